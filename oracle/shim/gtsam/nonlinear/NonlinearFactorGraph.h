#pragma once
#include "../../shim_gtsam.h"
