// oracle/shim: TEST INFRASTRUCTURE ONLY.
#pragma once
#include <pcl/point_types.h>
