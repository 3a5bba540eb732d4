// Microbenchmark: issue rate / throughput of tcgen05.mma shapes used by the attention kernels.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../../dualhyp_b200/csrc mma_rate.cu -o mma_rate -lcuda
#include <cstdio>
#include "common.cuh"
using namespace dh;
namespace dh { void set_error(const char*, ...) {} }

// mode 0: SS M128 N128 K16 ; 1: SS M128 N64 (B MN-major) ; 2: TS M128 N64 (B MN-major) ; 3: TS M128 N128 (B K-major) ; 4: SS M128 N256
template <int mode>
__global__ void __launch_bounds__(640, 1) rate_kernel(int reps, long long* out, int bg) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (warp == 0) { tmem_alloc(&slot, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  __shared__ volatile int stop;
  if (threadIdx.x == 0) stop = 0;
  __syncthreads();
  if (warp >= 4) {
    if (bg) {
      const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
      uint32_t v[32], w[16];
      for (int i = 0; i < 16; ++i) w[i] = i;
      while (!stop) {
        if (bg & 1) { tmem_ld32(tmem + lane_off + 384 + 32 * ((warp >> 2) & 3), v); tmem_wait_ld(); }
        if (bg & 2) { tmem_st16(tmem + lane_off + 320 + 16 * ((warp >> 2) & 3), w); tmem_wait_st(); }
        if (bg & 8) { float x0 = v[0], x1 = v[1], x2 = v[2], x3 = v[3], x4 = v[4], x5 = v[5], x6 = v[6], x7 = v[7];
#pragma unroll
          for (int i = 0; i < 64; ++i) { x0 = fmaf(x0, 1.0001f, 0.5f); x1 = fmaf(x1, 1.0001f, 0.5f); x2 = fmaf(x2, 1.0001f, 0.5f); x3 = fmaf(x3, 1.0001f, 0.5f); x4 = fmaf(x4, 1.0001f, 0.5f); x5 = fmaf(x5, 1.0001f, 0.5f); x6 = fmaf(x6, 1.0001f, 0.5f); x7 = fmaf(x7, 1.0001f, 0.5f); }
          w[1] = __float_as_uint(x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7); }
        if (bg & 4) { float x = v[0]; for (int i = 0; i < 32; ++i) x = fast_exp2(x) + 1.f; w[0] = __float_as_uint(x); }
      }
    }
  }
  if (threadIdx.x == 0) {
    const uint32_t a = smem_u32(smem), b = smem_u32(smem + 32768);
    long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if (mode == 0) tc_mma_ss(tmem, umma_desc_sw128(a + (k & 3) * 32, 16, 1024), umma_desc_sw128(b + (k & 3) * 32, 16, 1024), umma_idesc_bf16(128, 128, 0, 0), 1);
        else if (mode == 1) tc_mma_ss(tmem, umma_desc_sw128(a + (k & 3) * 32, 16, 1024), umma_desc_sw128(b + k * 2048, 8192, 1024), umma_idesc_bf16(128, 64, 0, 1), 1);
        else if (mode == 2) tc_mma_ts(tmem, tmem + 256 + k * 8, umma_desc_sw128(b + k * 2048, 8192, 1024), umma_idesc_bf16(128, 64, 0, 1), 1);
        else if (mode == 3) tc_mma_ts(tmem, tmem + 256 + k * 8, umma_desc_sw128(b + (k & 3) * 32, 16, 1024), umma_idesc_bf16(128, 128, 0, 0), 1);
        else if (mode == 4) tc_mma_ss(tmem, umma_desc_sw128(a + (k & 3) * 32, 16, 1024), umma_desc_sw128(b + (k & 3) * 32, 16, 1024), umma_idesc_bf16(128, 256, 0, 0), 1);
        else if (mode == 5) tc_mma_ts(tmem, tmem + 256 + k * 8, umma_desc_sw128(b + (k & 3) * 32, 16, 1024), umma_idesc_bf16(128, 64, 0, 0), 1);
        else if (mode == 6) tc_mma_ss(tmem, umma_desc_sw128(a + (k & 3) * 32, 16, 1024), umma_desc_sw128(b + (k & 3) * 32, 16, 1024), umma_idesc_bf16(128, 64, 0, 0), 1);
      }
    }
    long long t1 = clock64();
    tc_commit(&bar);
    mbar_wait(&bar, 0);
    long long t2 = clock64();
    if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
    stop = 1;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

int main() {
  long long* d;
  cudaMalloc(&d, 16);
  cudaFuncSetAttribute(rate_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(rate_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(rate_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(rate_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(rate_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(rate_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(rate_kernel<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  const char* names[] = {"SS M128 N128 (K-major B)", "SS M128 N64 (MN-major B)", "TS M128 N64 (MN-major B)", "TS M128 N128 (K-major B)",
                         "SS M128 N256", "TS M128 N64 (K-major B)", "SS M128 N64 (K-major B)"};
  for (int bg : {0, 8, 15}) {
    const int grid = 148;
    for (int mode = 0; mode < 7; ++mode) {
      const int reps = 64;
      switch (mode) {
        case 0: rate_kernel<0><<<grid, 640, 100 * 1024>>>(reps, d, bg); break;
        case 1: rate_kernel<1><<<grid, 640, 100 * 1024>>>(reps, d, bg); break;
        case 2: rate_kernel<2><<<grid, 640, 100 * 1024>>>(reps, d, bg); break;
        case 3: rate_kernel<3><<<grid, 640, 100 * 1024>>>(reps, d, bg); break;
        case 4: rate_kernel<4><<<grid, 640, 100 * 1024>>>(reps, d, bg); break;
        case 5: rate_kernel<5><<<grid, 640, 100 * 1024>>>(reps, d, bg); break;
        default: rate_kernel<6><<<grid, 640, 100 * 1024>>>(reps, d, bg); break;
      }
      cudaError_t e = cudaDeviceSynchronize();
      long long h[2];
      cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
      printf("bg %d grid %3d  %-28s issue %6.1f clk/mma   complete %6.1f clk/mma   (%s)\n", bg, grid, names[mode], h[0] / (8.0 * reps), h[1] / (8.0 * reps),
             cudaGetErrorString(e));
    }
  }
  return 0;
}
