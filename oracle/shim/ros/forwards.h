// oracle/shim: TEST INFRASTRUCTURE ONLY. See ros/shim_bus.h.
#pragma once
#include <ros/shim_bus.h>
