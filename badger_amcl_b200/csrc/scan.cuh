// Small device-wide primitives shared by the resampler and the kd-tree builder.
#pragma once
#include "common.cuh"

namespace bamcl
{
// Exclusive prefix sum of n int32 values (in may alias out); *total_out (device int) receives the sum.
// scratch must hold scan_scratch_ints(n) ints.
size_t scan_scratch_ints(long n);
int exclusive_scan_i32(bamcl_ctx* ctx, const int* in, int* out, long n, int* scratch, int* total_out);
}  // namespace bamcl
