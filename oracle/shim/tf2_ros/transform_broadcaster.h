// oracle/shim: TEST INFRASTRUCTURE ONLY. Nothing of tf2_ros is used by the hot path.
#pragma once
