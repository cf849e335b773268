#pragma once
#include "foundation/hg_math_types.h"
