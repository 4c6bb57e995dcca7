// qp_common.cuh - device-side parameter block, batch pointers and small helpers for the LPV-MPC QP kernels.
//
// Replaces (reference paths relative to planner/lib/plan_lib/):
//   distributedPlanner/LPV_Planner.py:138-157 (planes, weights, linearisation, QP data) and the OSQP call :222-249.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "qp_types.h"

#define QP_THREADS 128
#define QP_WARPS (QP_THREADS / 32)

// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Block-wide reductions over QP_THREADS threads; `red` is a QP_WARPS-entry shared scratch. All threads get the result.
__device__ __forceinline__ double block_max(double v, double* red) {
  v = warp_max(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double r = red[0];
#pragma unroll
  for (int w = 1; w < QP_WARPS; ++w) r = fmax(r, red[w]);
  return r;
}
__device__ __forceinline__ double block_min(double v, double* red) { return -block_max(-v, red); }
__device__ __forceinline__ double block_sum(double v, double* red) {
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double r = red[0];
#pragma unroll
  for (int w = 1; w < QP_WARPS; ++w) r += red[w];
  return r;
}
__device__ __forceinline__ int block_or(int v, int* redi) {
  v = __any_sync(0xffffffffu, v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) redi[threadIdx.x >> 5] = v;
  __syncthreads();
  int r = 0;
#pragma unroll
  for (int w = 0; w < QP_WARPS; ++w) r |= redi[w];
  return r;
}

// Track lookup with the semantics of utilities/misc.py:78-101 / :105-126: wrap while s > L, clamp s < 0,
// segment i with s0[i] <= s < s0[i]+len[i]; -1 where the reference's int(np.where(...)) raises.
// The wrap is one floor-based reduction plus a fix-up that keeps the reference's s == k L behaviour (its loop stops at
// the first value <= L, so an exact multiple of L ends at L, where no segment matches); |s| beyond 1e6 track lengths
// is reported as a failed lookup instead of being walked down.
__device__ __forceinline__ int map_segment(const DevParams& P, double s) {
  const double L = P.seg_s0[P.nseg - 1] + P.seg_len[P.nseg - 1];
  if (!(s == s) || !(L > 0.0) || fabs(s) > 1e6 * L) return -1;
  if (s > L) {
    s -= L * floor(s / L);
    if (s <= 0.0) s += L;
    while (s > L) s -= L;
  }
  if (s < 0.0) s = 0.0;
  int found = -1, cnt = 0;
  for (int i = 0; i < P.nseg; ++i)
    if (s >= P.seg_s0[i] && s < P.seg_s0[i] + P.seg_len[i]) { found = i; ++cnt; }
  return cnt == 1 ? found : -1;
}
