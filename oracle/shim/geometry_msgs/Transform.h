// oracle/shim: TEST INFRASTRUCTURE ONLY. See ros/shim_msgs.h.
#pragma once
#include <ros/shim_msgs.h>
