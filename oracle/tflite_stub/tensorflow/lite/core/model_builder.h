// Declaration-only stand-in (see c/c_api_types.h in this stub tree).
#pragma once
#include "tensorflow/lite/core/interpreter.h"
