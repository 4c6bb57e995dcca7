// peaks.cu - measured on-chip roofs of this B200 for the LPV-MPC solve kernel ("the FP64 peak is not in
// MEASURED_PEAKS.json - measure it on the box"): FP64 FMA throughput and shared-memory load bandwidth.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/ubench/peaks scripts/ubench/peaks.cu && scripts/ubench/peaks
// Prints one JSON line.  CUDA events, best of 5 after a warm-up launch.
#include <cstdio>
#include <cuda_runtime.h>

// 8 independent DFMA chains per thread: 8 warps x 8 chains per scheduler hide the ~9-cycle DFMA latency many times over
__global__ void __launch_bounds__(256) fma64(double* out, int iters, double a, double b) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
#pragma unroll 1
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
      x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}

// every thread streams 16-byte words out of a 32 KB shared buffer, conflict-free (consecutive lanes, consecutive words)
__global__ void __launch_bounds__(256) lds128(double* out, int iters) {
  __shared__ double2 buf[2048];
  for (int i = threadIdx.x; i < 2048; i += blockDim.x) buf[i] = make_double2(i, 1.0);
  __syncthreads();
  double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
#pragma unroll 1
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const double2 v = buf[(threadIdx.x + 256 * r + 8 * i) & 2047];
      if (r & 1) { s2 += v.x; s3 += v.y; } else { s0 += v.x; s1 += v.y; }
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = (s0 + s1) + (s2 + s3);
}

template <class F>
static float best_ms(F launch) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  launch();
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int r = 0; r < 5; ++r) {
    cudaEventRecord(e0);
    launch();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  return best;
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  const int blocks = p.multiProcessorCount * 8, threads = 256;
  double* out;
  cudaMalloc(&out, sizeof(double) * blocks * threads);
  const int it_f = 4096, it_l = 4096;
  const float ms_f = best_ms([&] { fma64<<<blocks, threads>>>(out, it_f, 1.0000001, 1e-9); });
  const float ms_l = best_ms([&] { lds128<<<blocks, threads>>>(out, it_l); });
  const double fmas = (double)blocks * threads * it_f * 64.0;
  const double bytes = (double)blocks * threads * it_l * 8.0 * 16.0;
  int clk = 0;
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"sm_clock_khz_max\": %d, \"fp64_fma_tflops\": %.3f, \"fp64_fma_per_clk_per_sm\": %.2f, "
         "\"smem_ld_gbs\": %.1f, \"smem_ld_bytes_per_clk_per_sm\": %.1f, \"how\": \"scripts/ubench/peaks.cu: 8 independent DFMA chains x 256 threads x 8 CTAs/SM "
         "(2 flop per FMA); LDS.128 conflict-free from 32 KB; CUDA events, best of 5\"}\n",
         p.name, p.multiProcessorCount, clk, 2.0 * fmas / (ms_f * 1e-3) / 1e12, fmas / (ms_f * 1e-3) / (clk * 1e3) / p.multiProcessorCount,
         bytes / (ms_l * 1e-3) / 1e9, bytes / (ms_l * 1e-3) / (clk * 1e3) / p.multiProcessorCount);
  cudaFree(out);
  return 0;
}
