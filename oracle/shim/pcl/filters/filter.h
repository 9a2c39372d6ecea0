#pragma once
#include "../../shim_pcl.h"
