// oracle/shim: TEST INFRASTRUCTURE ONLY. See octomap/OcTree.h.
#pragma once
#include <octomap/OcTree.h>
