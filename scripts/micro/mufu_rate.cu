// Microbenchmark: MUFU.EX2 throughput per SM (and an FMA-pipe polynomial exp2 for comparison).
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ float ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
// Cody-Waite + degree-3 minimax polynomial on the FMA pipe (valid for x <= 0 ... not too negative)
__device__ __forceinline__ float ex2_poly(float x) {
  x = fmaxf(x, -127.f);
  const float fl = floorf(x);
  const float f = x - fl;
  float pz = fmaf(f, 0.0555054f, 0.2402265f);
  pz = fmaf(pz, f, 0.6931472f);
  pz = fmaf(pz, f, 1.0f);
  return __int_as_float(__float_as_int(pz) + ((int)fl << 23));
}
template <int MODE>
__global__ void k(float* out, int iters) {
  float v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = -0.001f * (threadIdx.x + i);
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = (MODE == 0 ? ex2(v[i]) : ex2_poly(v[i])) - 1.5f;
  }
  long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) reinterpret_cast<long long*>(out)[100000] = t1 - t0;
}
int main() {
  float* d; cudaMalloc(&d, 4 << 20);
  for (int warps : {4, 8, 16, 32}) {
    for (int mode = 0; mode < 2; ++mode) {
      const int iters = 1000;
      if (mode == 0) k<0><<<148, warps * 32>>>(d, iters); else k<1><<<148, warps * 32>>>(d, iters);
      cudaDeviceSynchronize();
      long long c; cudaMemcpy(&c, reinterpret_cast<long long*>(d) + 100000, 8, cudaMemcpyDeviceToHost);
      const double per_sm = (double)warps * 32 * 16 * iters / c;
      printf("%s warps/SM %2d: %.1f results/clk/SM\n", mode == 0 ? "MUFU.EX2 " : "poly exp2", warps, per_sm);
    }
  }
}
