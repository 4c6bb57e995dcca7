// Development aid: dependent-chain latencies of the primitives the factorisation is made of (one warp, clock64).
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(double* out, long long* cyc, double seed) {
  __shared__ double sm[64];
  sm[threadIdx.x] = seed + threadIdx.x; sm[32 + threadIdx.x] = 1.0 / (seed + threadIdx.x);
  __syncwarp();
  double a = seed, b = 1.000001, c = 1e-7;
  long long t0, t1;
  const int R = 256;
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < R; ++i) a = fma(a, b, c);
  t1 = clock64(); if (threadIdx.x == 0) cyc[0] = (t1 - t0); out[0] = a;
  a = seed + 2.0;
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < R; ++i) a = rsqrt(a) + 3.0;
  t1 = clock64(); if (threadIdx.x == 0) cyc[1] = (t1 - t0); out[1] = a;
  a = seed + 2.0;
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < R; ++i) a = 7.0 / a + 1.0;
  t1 = clock64(); if (threadIdx.x == 0) cyc[2] = (t1 - t0); out[2] = a;
  a = seed + threadIdx.x;
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < R; ++i) a = __shfl_sync(0xffffffffu, a, (threadIdx.x + 1) & 31) + 1.0;
  t1 = clock64(); if (threadIdx.x == 0) cyc[3] = (t1 - t0); out[3] = a;
  int idx = threadIdx.x;
  a = 0;
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < R; ++i) { a += sm[idx & 63]; idx = (int)a & 31; }
  t1 = clock64(); if (threadIdx.x == 0) cyc[4] = (t1 - t0); out[4] = a;
  a = seed + 2.0;
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < R; ++i) { float f = rsqrtf((float)a); double y = (double)f; y = y * fma(-0.5 * a * y, y, 1.5); y = y * fma(-0.5 * a * y, y, 1.5); a = y + 3.0; }
  t1 = clock64(); if (threadIdx.x == 0) cyc[5] = (t1 - t0); out[5] = a;
  a = seed + 2.0;
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < R; ++i) a = sqrt(a) + 3.0;
  t1 = clock64(); if (threadIdx.x == 0) cyc[6] = (t1 - t0); out[6] = a;
  t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < R; ++i) __syncthreads();
  t1 = clock64(); if (threadIdx.x == 0) cyc[7] = (t1 - t0);
  a = seed + 2.0;
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < R; ++i) a = a * b;
  t1 = clock64(); if (threadIdx.x == 0) cyc[8] = (t1 - t0); out[8] = a;
  out[9] = sm[5];
}
int main() {
  double* o; long long* c; cudaMalloc(&o, 80); cudaMalloc(&c, 80);
  const char* nm[] = {"dfma chain", "rsqrt(double)+add", "div+add", "shfl(double)+add", "lds(dependent)+add+cvt", "rsqrtf seed + 2 NR +add", "sqrt+add", "__syncthreads (threads as launched)", "dmul chain"};
  for (int threads : {32, 128}) {
    k<<<1, threads>>>(o, c, 1.5); cudaDeviceSynchronize();
    long long h[10]; cudaMemcpy(h, c, 80, cudaMemcpyDeviceToHost);
    printf("threads %d\n", threads);
    for (int i = 0; i < 9; ++i) printf("  %-38s %7.1f cycles per step\n", nm[i], h[i] / 256.0);
  }
  return 0;
}
