#pragma once
#include "engine/scene.h"
