// hg::Scene facade — the reference's logging convention for this path: warn() and return a neutral value
// (harfang/foundation/log.h:7-20, used throughout engine/scene.cpp, e.g. :750-766).
#pragma once
namespace hg {
enum LogLevel { LL_Normal = 0x1, LL_Warning = 0x2, LL_Error = 0x4, LL_Debug = 0x8, LL_All = 0xff };
void set_log_level(int mask);
void set_log_hook(void (*on_log)(const char *msg, int mask, const char *details, void *user), void *user);
void log(const char *msg);
void warn(const char *msg);
void error(const char *msg);
} // namespace hg
