#pragma once
#include "../shim_ros.h"
