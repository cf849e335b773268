#pragma once
#include "foundation/generational_vector_list.h"
