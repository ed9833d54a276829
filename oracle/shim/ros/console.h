// oracle/shim: TEST INFRASTRUCTURE ONLY. ROS logging macros compiled to nothing.
#pragma once
#include <ros/assert.h>
#define ROS_DEBUG(...) ((void)0)
#define ROS_INFO(...) ((void)0)
#define ROS_WARN(...) ((void)0)
#define ROS_ERROR(...) ((void)0)
#define ROS_DEBUG_STREAM(x) ((void)0)
#define ROS_INFO_STREAM(x) ((void)0)
#define ROS_WARN_STREAM(x) ((void)0)
#define ROS_ERROR_STREAM(x) ((void)0)
#define ROS_INFO_DELAYED_THROTTLE(...) ((void)0)
