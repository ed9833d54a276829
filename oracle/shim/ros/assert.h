// oracle/shim: TEST INFRASTRUCTURE ONLY. ROS_ASSERT is compiled out (release ROS builds do the same).
#pragma once
#define ROS_ASSERT(x) ((void)0)
