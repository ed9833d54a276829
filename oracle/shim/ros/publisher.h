#pragma once
#include <ros/node_handle.h>
