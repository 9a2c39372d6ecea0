#pragma once
#include "../shim_cv.h"
