// Microbenchmark: how fast can ONE SM stream cold bytes from HBM into shared memory, and does the shape of the request
// matter?  Mode 0: cp.async.bulk of contiguous 16 KB chunks (what a tile-major pre-swizzled weight layout would allow).
// Mode 1: 128 separate 128-byte rows per chunk, rows `pitch` bytes apart (what a 128 x 64 bf16 TMA box over a K-major
// [N, K] weight does), issued as 128 small bulk copies.  N CTAs (one per SM), each with its own region, 8 chunks in
// flight.  Prints GB/s per CTA and in total.   nvcc -arch=sm_100a -O3 -o stream_rate stream_rate.cu && ./stream_rate
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(c)); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
  uint32_t ok = 0;
  while (!ok)
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes),
               "r"(smem_u32(bar)) : "memory");
}

constexpr int CHUNK = 16384, STAGES = 8;

template <int MODE>
__global__ void __launch_bounds__(128, 1) stream_kernel(const uint8_t* base, long long region_bytes, int chunks, long long pitch, long long* cycles) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t full[STAGES];
  const uint8_t* src = base + (long long)blockIdx.x * region_bytes;
  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) mbar_init(&full[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const long long t0 = clock64();
    for (int c = 0; c < chunks + STAGES; ++c) {
      const int s = c % STAGES;
      if (c >= STAGES) mbar_wait(&full[s], ((c / STAGES) - 1) & 1);      // chunk c - STAGES has landed: the slot is free again
      if (c < chunks) {
        mbar_expect(&full[s], CHUNK);
        if (MODE == 0) {
          bulk_g2s(smem + s * CHUNK, src + (long long)c * CHUNK, CHUNK, &full[s]);
        } else {
          // chunk c = k-block c of a [128 rows, K] slab: row r at src + r * pitch + c * 128
          for (int r = 0; r < 128; ++r) bulk_g2s(smem + s * CHUNK + r * 128, src + r * pitch + (long long)c * 128, 128, &full[s]);
        }
      }
    }
    cycles[blockIdx.x] = clock64() - t0;
  }
}

int main() {
  const long long total = 6LL << 30;
  uint8_t* d;
  cudaMalloc(&d, total);
  cudaMemset(d, 1, total);
  long long* cyc;
  cudaMalloc(&cyc, 148 * 8);
  cudaFuncSetAttribute(stream_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, CHUNK * STAGES);
  cudaFuncSetAttribute(stream_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, CHUNK * STAGES);
  int mhz = 0;
  cudaDeviceGetAttribute(&mhz, cudaDevAttrClockRate, 0);
  for (int mode = 0; mode < 3; ++mode)
    for (int n : {1, 16, 88, 148}) {
      const int chunks = mode == 2 ? 32 : 512;              // 8 MB per CTA, or the 512 KB a decode GEMM tile streams
      const long long pitch = 4096;                         // K = 2048 bf16 rows
      const long long region = (long long)chunks * CHUNK;
      cudaEvent_t e0, e1;
      cudaEventCreate(&e0); cudaEventCreate(&e1);
      float best = 1e9;
      for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        if (mode != 1) stream_kernel<0><<<n, 128, CHUNK * STAGES>>>(d + (long long)rep * (1LL << 31), region, chunks, pitch, cyc);
        else stream_kernel<1><<<n, 128, CHUNK * STAGES>>>(d + (long long)rep * (1LL << 31), 128 * pitch, 32, pitch, cyc);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
      }
      const double bytes = mode != 1 ? (double)chunks * CHUNK : 32.0 * CHUNK;
      printf("mode %d (%s) CTAs %3d: %.1f us, %.1f GB/s per CTA, %.2f TB/s total  [%s]\n", mode, mode == 0 ? "contiguous 16 KB x 512" : mode == 1 ? "128 rows x 128 B x 32" : "contiguous 16 KB x 32",
             n, best * 1e3, bytes / (best * 1e-3) / 1e9, bytes * n / (best * 1e-3) / 1e12, cudaGetErrorString(cudaGetLastError()));
    }
  return 0;
}
