// Persistent, warp-specialised bf16 GEMM for sm_100a: TMA -> swizzled smem ring -> tcgen05.mma (single
// issuing thread) -> double-buffered fp32 accumulators in TMEM -> epilogue warps (tcgen05.ld) with the
// fused epilogues the DualHyp decoder needs (RoPE on the interleaved QKV row, SwiGLU fwd/bwd, residual
// add, LoRA delta via a second K source).  Replaces ger/model.py:213,266,313-316, ger/lora.py:162-166,
// 388-402 and the lm_head linear ger/lora.py:70-71.
//
// Roles (256 threads, 1 CTA / SM): warp 0 = TMA producer, warp 1 = MMA issuer, warp 2 = TMEM allocator,
// warps 4-7 = epilogue (warp w owns TMEM lanes 32*(w%4) .. +31, one accumulator row per thread).
#include <stdlib.h>

#include "common.cuh"
#include "../../include/dualhyp.h"

namespace dh {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int GEMM_THREADS = 256;

struct GemmParams {
  int M, N, K1, K2;
  int tiles_m, tiles_n;
  int a_mn, b_mn, a2_mn, b2_mn;
  int accumulate;
  float alpha;
  int seq_len, q_per_kv;
  void* out; long long ldo;
  void* out2; long long ldo2;
  const void* aux; long long ldaux;
  const float* col_scale;
  const float* col_bias;
  const bf16* rope_cos;
  const bf16* rope_sin;
  const int* positions;
  const int* m_limit;      // optional device-side row count: tiles at or beyond it are skipped
  int splits;              // gemm_kernel: K split into `splits` ranges, range s -> output slice s (fp32 partial products)
  long long split_stride;  // elements between output slices
  int sk_stages;           // skinny kernel: depth of the TMA ring
  int sk_partial;          // skinny kernel: gridDim.x CTAs per tile each store their K range's product to slice blockIdx.x of out
  int b_frozen;            // skinny kernel: B (first source) is a frozen weight -- its tiles may be fetched before pdl_wait()
  int band_m;              // pair kernel: M-tiles per band of the tile order (see pair_tile_coords)
};

// Tile order of the persistent CTA-pair kernel.  Tiles are walked band by band: a band is `band_m` consecutive 256-row M-tiles
// crossed with ALL N-tiles, M fastest inside the band.  The 74 pairs running at one time then share a few weight tiles, and
// the band's A rows (band_m x 256 x K x 2 bytes, sized by the launcher to a fraction of the 126 MB L2) stay L2-resident while
// the band sweeps N -- A and W are each read from HBM about once.  With band_m >= tiles_m (every training shape of the bench:
// A is 27 MB) this is the plain M-fastest order.  Without bands a prefill GEMM over 100 k rows re-streamed its 500 MB A operand
// once per N-tile: 22 GB of DRAM traffic per launch, 1105 TF/s.
__device__ __forceinline__ void pair_tile_coords(int tile, int tiles_m, int tiles_n, int band_m, int& tm, int& tn) {
  const int per_band = band_m * tiles_n;
  const int b = tile / per_band, r = tile - b * per_band;
  const int m_first = b * band_m;
  const int rows = tiles_m - m_first < band_m ? tiles_m - m_first : band_m;     // the last band may be short
  tn = r / rows;
  tm = m_first + (r - tn * rows);
}

template <int BN> struct GemmCfg {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = (BN == 256) ? 4 : (BN == 128 ? 6 : 8);
  static constexpr int TMEM_COLS = (2 * BN <= 32) ? 32 : (2 * BN <= 64 ? 64 : (2 * BN <= 128 ? 128 : (2 * BN <= 256 ? 256 : 512)));
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
};

// ------------------------------------------------------------------------------------------------
// epilogues: one thread owns `row`, v[0..63] are 64 consecutive accumulator columns starting at col0
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void store_bf16x64(bf16* dst, const float (&f)[64]) {
  uint4* d4 = reinterpret_cast<uint4*>(dst);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    uint4 u;
    u.x = pack_bf16x2(f[8 * i + 0], f[8 * i + 1]);
    u.y = pack_bf16x2(f[8 * i + 2], f[8 * i + 3]);
    u.z = pack_bf16x2(f[8 * i + 4], f[8 * i + 5]);
    u.w = pack_bf16x2(f[8 * i + 6], f[8 * i + 7]);
    d4[i] = u;
  }
}
__device__ __forceinline__ void load_bf16x32(const bf16* src, float (&f)[32]) {
  const uint4* s4 = reinterpret_cast<const uint4*>(src);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    uint4 u = s4[i];
    float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
    f[8 * i + 0] = a.x; f[8 * i + 1] = a.y; f[8 * i + 2] = b.x; f[8 * i + 3] = b.y;
    f[8 * i + 4] = c.x; f[8 * i + 5] = c.y; f[8 * i + 6] = d.x; f[8 * i + 7] = d.y;
  }
}

__device__ __forceinline__ void unpack_bf16x8(const uint4& u, float (&f)[8]) {
  float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 pack_bf16x8(const float (&f)[8]) {
  uint4 u;
  u.x = pack_bf16x2(f[0], f[1]); u.y = pack_bf16x2(f[2], f[3]); u.z = pack_bf16x2(f[4], f[5]); u.w = pack_bf16x2(f[6], f[7]);
  return u;
}

// sigmoid(x) = 0.5 + 0.5 tanh(x / 2): ONE special-function op (MUFU.TANH) instead of the two of 1 / (1 + exp(-x)) (EX2 + RCP).  The
// SwiGLU-backward epilogue needs a sigmoid for every accumulator element: 64 K MUFU ops per CTA and tile at 16 per clock were as
// long as the K = 2048 main loop (4096 clocks); tanh.approx is accurate to ~2^-11, the outputs are rounded to bf16 (2^-9).
__device__ __forceinline__ float fast_sigmoid(float x) {
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(0.5f * x));
  return fmaf(0.5f, t, 0.5f);
}

template <int EPI>
__device__ __forceinline__ void epilogue_chunk(const GemmParams& p, int row, int col0, float (&v)[64]) {
  if constexpr (EPI == DH_EPI_BF16) {
#pragma unroll
    for (int j = 0; j < 64; ++j) v[j] *= p.alpha;
    if (p.col_bias) {
#pragma unroll
      for (int j = 0; j < 64; ++j) v[j] = bf16_round(v[j]) + __ldg(p.col_bias + col0 + j);
    }
    if (p.col_scale) {
#pragma unroll
      for (int j = 0; j < 64; ++j) v[j] *= __ldg(p.col_scale + col0 + j);
    }
    store_bf16x64(reinterpret_cast<bf16*>(p.out) + (long long)row * p.ldo + col0, v);
  } else if constexpr (EPI == DH_EPI_F32) {
    float4* o4 = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + (long long)row * p.ldo + col0);
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      float4 r = make_float4(v[4 * i] * p.alpha, v[4 * i + 1] * p.alpha, v[4 * i + 2] * p.alpha, v[4 * i + 3] * p.alpha);
      if (p.accumulate) {
        float4 o = o4[i];
        r.x += o.x; r.y += o.y; r.z += o.z; r.w += o.w;
      }
      o4[i] = r;
    }
  } else if constexpr (EPI == DH_EPI_RESID_F32) {
    const float4* r4 = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.aux) + (long long)row * p.ldaux + col0);
    float4* o4 = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + (long long)row * p.ldo + col0);
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      float4 r = r4[i];
      r.x += bf16_round(v[4 * i]); r.y += bf16_round(v[4 * i + 1]);
      r.z += bf16_round(v[4 * i + 2]); r.w += bf16_round(v[4 * i + 3]);
      o4[i] = r;
    }
  } else if constexpr (EPI == DH_EPI_RESID_BF16) {
    const bf16* r = reinterpret_cast<const bf16*>(p.aux) + (long long)row * p.ldaux + col0;
    float a[32];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      load_bf16x32(r + 32 * h, a);
#pragma unroll
      for (int j = 0; j < 32; ++j) v[32 * h + j] = a[j] + bf16_round(v[32 * h + j]);
    }
    store_bf16x64(reinterpret_cast<bf16*>(p.out) + (long long)row * p.ldo + col0, v);
  } else if constexpr (EPI == DH_EPI_QKV_ROPE) {
    const int slot = (col0 >> 6) % (p.q_per_kv + 2);
    if (slot <= p.q_per_kv) {  // q heads and the k head rotate; v passes through
      const int pos = p.positions ? __ldg(p.positions + row) : (row % p.seq_len);
      float c[32], s[32];
      load_bf16x32(p.rope_cos + (long long)pos * 64, c);   // table halves are identical (repeat(1,2))
      load_bf16x32(p.rope_sin + (long long)pos * 64, s);
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        float x1 = bf16_round(v[j]), x2 = bf16_round(v[j + 32]);
        v[j] = x1 * c[j] - x2 * s[j];
        v[j + 32] = x2 * c[j] + x1 * s[j];
      }
    }
    store_bf16x64(reinterpret_cast<bf16*>(p.out) + (long long)row * p.ldo + col0, v);
  } else if constexpr (EPI == DH_EPI_SWIGLU) {
    if (p.out) store_bf16x64(reinterpret_cast<bf16*>(p.out) + (long long)row * p.ldo + col0, v);
    uint4* d4 = reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(p.out2) + (long long)row * p.ldo2 + (col0 >> 1));
    float a[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      float h1 = bf16_round(v[j]), h2 = bf16_round(v[j + 32]);
      a[j] = h1 * fast_sigmoid(h1) * h2;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint4 u;
      u.x = pack_bf16x2(a[8 * i + 0], a[8 * i + 1]); u.y = pack_bf16x2(a[8 * i + 2], a[8 * i + 3]);
      u.z = pack_bf16x2(a[8 * i + 4], a[8 * i + 5]); u.w = pack_bf16x2(a[8 * i + 6], a[8 * i + 7]);
      d4[i] = u;
    }
  } else if constexpr (EPI == DH_EPI_SWIGLU_BWD) {
    // acc columns col0..col0+63 of d(act); h12 / d(h12) live in 64-wide [h1 x32 | h2 x32] blocks.
    const bf16* h = reinterpret_cast<const bf16*>(p.aux) + (long long)row * p.ldaux + 2 * col0;
    bf16* o = reinterpret_cast<bf16*>(p.out) + (long long)row * p.ldo + 2 * col0;
#pragma unroll
    for (int g = 0; g < 2; ++g) {
      float h1[32], h2[32], d[64];
      load_bf16x32(h + 64 * g, h1);
      load_bf16x32(h + 64 * g + 32, h2);
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        float da = bf16_round(v[32 * g + j]);
        float sig = fast_sigmoid(h1[j]);
        float silu = h1[j] * sig;
        d[j] = da * h2[j] * (sig * (1.f + h1[j] * (1.f - sig)));
        d[j + 32] = da * silu;
      }
      store_bf16x64(o + 64 * g, d);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// kernel
// ------------------------------------------------------------------------------------------------
template <int BN, int EPI>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2, const GemmParams p) {
  using Cfg = GemmCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES);
  uint64_t* full_bar = bars;                 // [STAGES]  TMA -> MMA
  uint64_t* empty_bar = bars + STAGES;       // [STAGES]  MMA -> TMA
  uint64_t* tfull_bar = bars + 2 * STAGES;   // [2]       MMA -> epilogue
  uint64_t* tempty_bar = bars + 2 * STAGES + 2;  // [2]   epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // shfl: a warp-uniform value for the compiler
  const int lane = threadIdx.x & 31;
  const int nkb1 = (p.K1 + BK - 1) / BK;   // a ragged K tail is zero-filled by TMA
  const int nkb = nkb1 + (p.K2 + BK - 1) / BK;
  pdl_launch_dependents();

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (p.K2 > 0) { tma_prefetch_desc(&tmA2); tma_prefetch_desc(&tmB2); }
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 128); }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  pdl_wait();          // the prologue above overlapped the previous kernel's tail; global memory is read from here on
  int tiles_m = p.tiles_m;
  if (p.m_limit) { const int t = (__ldg(p.m_limit) + BM - 1) / BM; tiles_m = t < tiles_m ? t : tiles_m; }
  const int real_tiles = tiles_m * p.tiles_n;
  const int num_tiles = real_tiles * p.splits;   // virtual tiles: (output tile, K range); range s writes output slice s

  if (warp == 0 && lane == 0) {
    // ===================== TMA producer =====================
    uint32_t it = 0;
    for (int vt = blockIdx.x; vt < num_tiles; vt += gridDim.x) {
      const int tile = vt % real_tiles, split = vt / real_tiles;
      const int m0 = (tile % tiles_m) * BM;
      const int n0 = (tile / tiles_m) * BN;
      const int kb0 = (int)((long long)nkb * split / p.splits), kb1 = (int)((long long)nkb * (split + 1) / p.splits);
      for (int kb = kb0; kb < kb1; ++kb, ++it) {
        const int s = it % STAGES;
        mbar_wait(&empty_bar[s], ((it / STAGES) & 1) ^ 1);
        uint8_t* sa = smem + s * Cfg::STAGE_BYTES;
        uint8_t* sb = sa + Cfg::A_BYTES;
        mbar_expect_tx(&full_bar[s], Cfg::STAGE_BYTES);
        const bool second = kb >= nkb1;
        const int k0 = (second ? kb - nkb1 : kb) * BK;
        const CUtensorMap* ma = second ? &tmA2 : &tmA;
        const CUtensorMap* mb = second ? &tmB2 : &tmB;
        if (!(second ? p.a2_mn : p.a_mn)) {
          tma_load_2d(sa, ma, k0, m0, &full_bar[s]);                       // box {64 k, 128 m}
        } else {
#pragma unroll
          for (int j = 0; j < BM / 64; ++j) tma_load_2d(sa + j * 8192, ma, m0 + 64 * j, k0, &full_bar[s]);  // box {64 m, 64 k}
        }
        if (!(second ? p.b2_mn : p.b_mn)) {
          tma_load_2d(sb, mb, k0, n0, &full_bar[s]);                       // box {64 k, BN n}
        } else {
#pragma unroll
          for (int j = 0; j < BN / 64; ++j) tma_load_2d(sb + j * 8192, mb, n0 + 64 * j, k0, &full_bar[s]);  // box {64 n, 64 k}
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer: the whole warp walks the loop (uniform operands), one elected lane issues =====================
    uint32_t it = 0, tcount = 0;
    for (int vt = blockIdx.x; vt < num_tiles; vt += gridDim.x, ++tcount) {
      const int split = vt / real_tiles;
      const int kb0 = (int)((long long)nkb * split / p.splits), kb1 = (int)((long long)nkb * (split + 1) / p.splits);
      const int acc = tcount & 1;
      mbar_wait(&tempty_bar[acc], ((tcount >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * BN;
      for (int kb = kb0; kb < kb1; ++kb, ++it) {
        const int s = it % STAGES;
        mbar_wait(&full_bar[s], (it / STAGES) & 1);
        tc_fence_after();
        const bool second = kb >= nkb1;
        const int a_mn = second ? p.a2_mn : p.a_mn;
        const int b_mn = second ? p.b2_mn : p.b_mn;
        const uint32_t idesc = umma_idesc_bf16(BM, BN, a_mn, b_mn);
        const uint32_t sa = smem_u32(smem + s * Cfg::STAGE_BYTES);
        const uint32_t sb = sa + Cfg::A_BYTES;
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            // K-major: 16 k-elements = 32 bytes inside the 128B swizzle row.  MN-major: 16 k-rows = two
            // 8-row atoms of 1024 B; 64-wide MN groups are 8 KB apart (LBO), 8-row atoms 1 KB apart (SBO).
            const uint64_t da = a_mn ? umma_desc_sw128(sa + k * 2048, 8192, 1024) : umma_desc_sw128(sa + k * 32, 16, 1024);
            const uint64_t db = b_mn ? umma_desc_sw128(sb + k * 2048, 8192, 1024) : umma_desc_sw128(sb + k * 32, 16, 1024);
            tc_mma_ss(d_tmem, da, db, idesc, ((kb - kb0) | k) != 0);
          }
          tc_commit(&empty_bar[s]);   // frees the smem slot when these MMAs have drained
          if (kb == kb1 - 1) tc_commit(&tfull_bar[acc]);   // accumulator complete
        }
        __syncwarp();
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    const int q = warp & 3;
    uint32_t tcount = 0;
    for (int vt = blockIdx.x; vt < num_tiles; vt += gridDim.x, ++tcount) {
      const int tile = vt % real_tiles, split = vt / real_tiles;
      const int acc = tcount & 1;
      const int m0 = (tile % tiles_m) * BM;
      const int n0 = (tile / tiles_m) * BN;
      GemmParams ps = p;
      if (EPI == DH_EPI_F32 && p.splits > 1) ps.out = reinterpret_cast<float*>(p.out) + (long long)split * p.split_stride;
      mbar_wait(&tfull_bar[acc], (tcount >> 1) & 1);
      tc_fence_after();
      const int row = m0 + q * 32 + lane;
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN;
#pragma unroll 1
      for (int c = 0; c < BN / 64; ++c) {
        uint32_t r0[32], r1[32];
        tmem_ld32(taddr + c * 64, r0);
        tmem_ld32(taddr + c * 64 + 32, r1);
        tmem_wait_ld();
        float v[64];
#pragma unroll
        for (int j = 0; j < 32; ++j) { v[j] = __uint_as_float(r0[j]); v[j + 32] = __uint_as_float(r1[j]); }
        if (row < p.M) epilogue_chunk<EPI>(ps, row, n0 + c * 64, v);
      }
      tc_fence_before();
      mbar_arrive(&tempty_bar[acc]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------
// Skinny GEMM (M <= 128: the decode step, one token per sequence) -- weight streaming, HBM-bound.
// A 128 x 128 output tile per CTA leaves only N / 128 tiles (16 when N = 2048), so K is split over a thread-block
// CLUSTER of S <= 4 CTAs per tile: each CTA streams its K range through an 8-stage TMA ring into tcgen05.mma,
// parks its fp32 partial tile in its own shared memory, and after one cluster barrier the tile's (row, 64-column)
// units are dealt round-robin to the S CTAs, each summing the S partials over DISTRIBUTED SHARED MEMORY in rank
// order (deterministic) and running the fused epilogue.  No global workspace, no atomics.
// 128 < M <= 256: two activation boxes per stage and two TMEM accumulators, the weight tile is streamed once for both.
// M <= 64 loads only a 64-row A box: the accumulator rows 64..127 are then garbage computed from whatever the
// B tile behind it holds (MMA rows are independent) and are never read, so a stage is 8 KB of A + 16 KB of W.
// Roles as gemm_kernel: warp 0 TMA, warp 1 MMA, warp 2 TMEM, warps 4-7 epilogue.
// ------------------------------------------------------------------------------------------------
constexpr int SK_BN = 128;
constexpr int SK_B_BYTES = SK_BN * BK * 2;
constexpr int SK_PAD = SK_BN + 4;               // fp32 partial row pitch (floats): 16-byte aligned, conflict-free rows
constexpr int SK_RING_BYTES = 192 * 1024;
constexpr int SK_MAX_STAGES = 8;

// cluster barrier without .aligned: the role branches above leave warp 0 / 1 possibly un-converged
__device__ __forceinline__ void cluster_sync_any() {
  asm volatile("barrier.cluster.arrive.release;\n\tbarrier.cluster.wait.acquire;" ::: "memory");
}
__device__ __forceinline__ float4 ld_dsmem_f4(uint32_t cluster_addr) {
  float4 v;
  asm("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(cluster_addr));
  return v;
}

template <int EPI>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_skinny_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                   const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2, const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int MB = p.M > BM ? 2 : 1;                       // row blocks: up to 256 activation rows stream the same weight tile
  const int a_box_bytes = p.M <= 64 ? 64 * BK * 2 : BM * BK * 2;
  const int a_bytes = a_box_bytes * MB;
  const int stage_bytes = a_bytes + SK_B_BYTES;
  const int stages = p.sk_stages;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + stages * stage_bytes);
  uint64_t* full_bar = bars;                       // [stages]
  uint64_t* empty_bar = bars + SK_MAX_STAGES;      // [stages]
  uint64_t* tfull_bar = bars + 2 * SK_MAX_STAGES;  // [1]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * SK_MAX_STAGES + 1);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const int KS = gridDim.x, split = blockIdx.x, tile = blockIdx.y;     // the cluster spans x: rank == split
  const int S = p.sk_partial ? 1 : KS;          // CTAs whose partial tiles are reduced here (partial mode: none, the consumer sums)
  const int n0 = tile * SK_BN;
  const int nkb1 = (p.K1 + BK - 1) / BK;
  const int nkb = nkb1 + (p.K2 + BK - 1) / BK;
  const int kb_begin = (int)((long long)nkb * split / KS), kb_end = (int)((long long)nkb * (split + 1) / KS);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (p.K2 > 0) { tma_prefetch_desc(&tmA2); tma_prefetch_desc(&tmB2); }
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < stages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    mbar_init(tfull_bar, 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 2 * SK_BN);                    // one accumulator per row block
    tmem_relinquish();
  }
  pdl_launch_dependents();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

  if (warp == 0 && lane == 0) {
    // A frozen weight does not depend on the previous kernel: its first ring of tiles is requested BEFORE pdl_wait(), so
    // the weight stream starts while the predecessor (a consumer on a few SMs, another GEMM's tail) is still running;
    // the activation boxes follow after the wait and complete the same stage barriers.
    int n_pre = 0;
    if (p.b_frozen) {
      const int first_end = kb_end < nkb1 ? kb_end : nkb1;
      for (int kb = kb_begin; kb < first_end && n_pre < stages; ++kb, ++n_pre) {
        mbar_expect_tx(&full_bar[n_pre], stage_bytes);
        tma_load_2d(smem + n_pre * stage_bytes + a_bytes, &tmB, kb * BK, n0, &full_bar[n_pre]);
      }
    }
    pdl_wait();
    uint32_t it = 0;
    for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
      const int s = it % stages;
      uint8_t* sa = smem + s * stage_bytes;
      uint8_t* sb = sa + a_bytes;
      const bool second = kb >= nkb1;
      const int k0 = (second ? kb - nkb1 : kb) * BK;
      if ((int)it >= n_pre) {
        mbar_wait(&empty_bar[s], ((it / stages) & 1) ^ 1);
        mbar_expect_tx(&full_bar[s], stage_bytes);
        tma_load_2d(sb, second ? &tmB2 : &tmB, k0, n0, &full_bar[s]);
      }
      tma_load_2d(sa, second ? &tmA2 : &tmA, k0, 0, &full_bar[s]);
      if (MB == 2) tma_load_2d(sa + a_box_bytes, second ? &tmA2 : &tmA, k0, BM, &full_bar[s]);
    }
  } else if (warp == 1) {
    pdl_wait();
    const uint32_t idesc = umma_idesc_bf16(BM, SK_BN, 0, 0);
    uint32_t it = 0;
    for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
      const int s = it % stages;
      mbar_wait(&full_bar[s], (it / stages) & 1);
      tc_fence_after();
      const uint32_t sa = smem_u32(smem + s * stage_bytes);
      const uint32_t sb = sa + a_bytes;
      if (elect_one()) {
        for (int rb = 0; rb < MB; ++rb) {
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)
            tc_mma_ss(tmem_base + rb * SK_BN, umma_desc_sw128(sa + rb * a_box_bytes + k * 32, 16, 1024), umma_desc_sw128(sb + k * 32, 16, 1024),
                      idesc, (it | k) != 0);
        }
        tc_commit(&empty_bar[s]);
        if (kb == kb_end - 1) tc_commit(tfull_bar);
      }
      __syncwarp();
    }
  } else if (warp >= 4 && S > 1) {
    pdl_wait();
    // park this CTA's partial rows in shared memory (the ring is idle once the last MMA has completed)
    const int q = warp & 3;
    mbar_wait(tfull_bar, 0);
    tc_fence_after();
    for (int rb = 0; rb < MB; ++rb) {
      const int r = rb * BM + q * 32 + lane;
      float* prow = reinterpret_cast<float*>(smem) + (size_t)r * SK_PAD;
#pragma unroll 1
      for (int c = 0; c < SK_BN / 32; ++c) {
        uint32_t r0[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + rb * SK_BN + c * 32, r0);
        tmem_wait_ld();
        if (r < p.M) {
#pragma unroll
          for (int i = 0; i < 8; ++i)
            *reinterpret_cast<float4*>(prow + c * 32 + 4 * i) =
                make_float4(__uint_as_float(r0[4 * i]), __uint_as_float(r0[4 * i + 1]), __uint_as_float(r0[4 * i + 2]), __uint_as_float(r0[4 * i + 3]));
        }
      }
    }
  }
  if (S > 1) {
    tc_fence_before();
    __syncthreads();
    cluster_sync_any();
  }
  if (warp >= 4) {
    pdl_wait();
    const int t = threadIdx.x - 128;
    if (S == 1 && EPI == DH_EPI_F32 && p.sk_partial) {
      // Slice `split` of the [KS][M][ldo] partial products.  A thread holds a ROW of the accumulator (tcgen05.ld 32x32b), so storing
      // from registers makes every store instruction touch 32 rows = 32 cache lines, 16 bytes each: 2048 line requests per warp
      // for a 256 x 128 tile.  The idle ring transposes instead (padded 32 x 36 tile per warp, conflict-free both ways): a store
      // instruction then writes 4 rows x 128 contiguous bytes.
      const int q = warp & 3;
      mbar_wait(tfull_bar, 0);
      tc_fence_after();
      float* outp = reinterpret_cast<float*>(p.out) + (long long)split * p.M * p.ldo + n0;
      const uint32_t stg = smem_u32(smem) + (uint32_t)q * (32 * 36 * 4);
      const int rsub = lane >> 3, cq = lane & 7;
#pragma unroll 1
      for (int rb = 0; rb < MB; ++rb) {
        const int row0 = rb * BM + q * 32;
        const int nrow = p.M - row0;                           // valid rows of this warp's 32-row block (may be <= 0)
#pragma unroll 1
        for (int g = 0; g < SK_BN / 32; ++g) {
          uint32_t rg[32];
          tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + rb * SK_BN + g * 32, rg);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 8; ++i)
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(stg + lane * 144 + i * 16), "f"(__uint_as_float(rg[4 * i]) * p.alpha),
                         "f"(__uint_as_float(rg[4 * i + 1]) * p.alpha), "f"(__uint_as_float(rg[4 * i + 2]) * p.alpha),
                         "f"(__uint_as_float(rg[4 * i + 3]) * p.alpha)
                         : "memory");
          __syncwarp();
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int rr = 4 * k + rsub;
            float4 v4;
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v4.x), "=f"(v4.y), "=f"(v4.z), "=f"(v4.w) : "r"(stg + rr * 144 + cq * 16));
            if (rr < nrow) *reinterpret_cast<float4*>(outp + (long long)(row0 + rr) * p.ldo + g * 32 + cq * 4) = v4;
          }
          __syncwarp();
        }
      }
    } else if (S == 1) {
      const int q = warp & 3;
      mbar_wait(tfull_bar, 0);
      tc_fence_after();
#pragma unroll 1
      for (int cc = 0; cc < MB * (SK_BN / 64); ++cc) {
        const int rb = cc / (SK_BN / 64), c = cc % (SK_BN / 64);
        const int row = rb * BM + q * 32 + lane;
        uint32_t r0[32], r1[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + rb * SK_BN + c * 64;
        tmem_ld32(taddr, r0);
        tmem_ld32(taddr + 32, r1);
        tmem_wait_ld();
        float v[64];
#pragma unroll
        for (int j = 0; j < 32; ++j) { v[j] = __uint_as_float(r0[j]); v[j + 32] = __uint_as_float(r1[j]); }
        if (row < p.M) {
          if (EPI == DH_EPI_F32 && p.sk_partial) {          // slice `split` of the [KS][M][ldo] partial products
            GemmParams ps = p;
            ps.out = reinterpret_cast<float*>(p.out) + (long long)split * p.M * p.ldo;
            epilogue_chunk<EPI>(ps, row, n0 + c * 64, v);
          } else {
            epilogue_chunk<EPI>(p, row, n0 + c * 64, v);
          }
        }
      }
    } else {
      // This CTA finishes the (row, 64-column) units u = split, split + S, ... of the tile.  Phase 1: all 128 epilogue
      // threads sum the S partials of those units over distributed shared memory, one float4 column group at a time
      // (S independent remote loads in flight per thread, ranks in fixed order => deterministic).  Phase 2: one thread
      // per unit runs the fused epilogue on the reduced 64 values.
      const int units = p.M * (SK_BN / 64);
      const int my_units = units > split ? (units - split + S - 1) / S : 0;
      float* red = reinterpret_cast<float*>(smem) + (size_t)p.M * SK_PAD;      // [my_units][64] behind the partial tile
      const uint32_t base = smem_u32(smem);
      for (int e = t; e < my_units * 16; e += 128) {
        const int j = e >> 4, i = e & 15;
        const int u = split + S * j;
        const uint32_t local = base + (uint32_t)(((u >> 1) * SK_PAD + (u & 1) * 64 + 4 * i) * 4);
        float4 x[8];
#pragma unroll
        for (int s2 = 0; s2 < 8; ++s2)
          if (s2 < S) x[s2] = ld_dsmem_f4(mapa_shared(local, (uint32_t)s2));
        float4 a = x[0];
#pragma unroll
        for (int s2 = 1; s2 < 8; ++s2)
          if (s2 < S) { a.x += x[s2].x; a.y += x[s2].y; a.z += x[s2].z; a.w += x[s2].w; }
        *reinterpret_cast<float4*>(red + j * 64 + 4 * i) = a;
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      for (int j = t; j < my_units; j += 128) {
        const int u = split + S * j;
        float v[64];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float4 a = *reinterpret_cast<const float4*>(red + j * 64 + 4 * i);
          v[4 * i] = a.x; v[4 * i + 1] = a.y; v[4 * i + 2] = a.z; v[4 * i + 3] = a.w;
        }
        epilogue_chunk<EPI>(p, u >> 1, n0 + (u & 1) * 64, v);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (S > 1) cluster_sync_any();       // nobody leaves while a peer may still read its partial tile
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * SK_BN);
  }
}

// ------------------------------------------------------------------------------------------------
// Stream-K decode GEMM (M <= 256, dh_gemm_args.workspace given): for GEMMs whose N / 128 tiles do not fill the machine and whose
// epilogue is tile-local (SwiGLU: 88 tiles on 148 SMs left 60 SMs idle and took 33 us at 256 rows).  The (128-column tile,
// 64-wide k-block) units are dealt out contiguously over ALL SMs, so every SM streams the same number of weight bytes; a CTA's
// run touches one or two tiles, and each (CTA, tile) segment leaves an fp32 partial tile in the workspace (TMEM -> a shared-memory
// transpose -> 128-byte row stores), then bumps the tile's arrival counter.  The CTA whose arrival completes a tile sums the
// tile's slices in CTA order (deterministic whoever arrives last) and runs the fused epilogue.  Nobody ever WAITS for another
// CTA, so there is no residency requirement and no deadlock to reason about; the counters are reset by the finisher.
// Roles as gemm_skinny_kernel: warp 0 TMA, warp 1 MMA, warp 2 TMEM, warps 4-7 drain / finish.
// ------------------------------------------------------------------------------------------------
constexpr int SKK_STG_BYTES = 4 * 32 * 32 * 4;      // one [32][32] fp32 transpose tile per epilogue warp

__host__ __device__ __forceinline__ int skk_unit_lo(int units, int c, int P) { return (int)((long long)units * c / P); }
__host__ __device__ __forceinline__ int skk_unit_owner(int units, int u, int P) { return (int)((((long long)u + 1) * P - 1) / units); }

template <int EPI>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_streamk_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmParams p,
                    float* __restrict__ ws, unsigned int* __restrict__ counters, int maxseg) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int MB = p.M > BM ? 2 : 1;
  const int a_box_bytes = p.M <= 64 ? 64 * BK * 2 : BM * BK * 2;
  const int a_bytes = a_box_bytes * MB;
  const int stage_bytes = a_bytes + SK_B_BYTES;
  const int stages = p.sk_stages;
  float* stg = reinterpret_cast<float*>(smem + stages * stage_bytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + stages * stage_bytes + SKK_STG_BYTES);
  uint64_t* full_bar = bars;                       // [stages]
  uint64_t* empty_bar = bars + SK_MAX_STAGES;      // [stages]
  uint64_t* tfull_bar = bars + 2 * SK_MAX_STAGES;  // [2]
  uint64_t* tempty_bar = tfull_bar + 2;            // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  int* last_flag = reinterpret_cast<int*>(tmem_slot + 1);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const int cta = blockIdx.x;
  const int nkb = p.K1 / BK;
  const int units = p.tiles_n * nkb;
  const int P = units < (int)gridDim.x ? units : (int)gridDim.x;   // CTAs that share the work: each owns at least one unit (the rest idle)
  const int u0 = cta < P ? skk_unit_lo(units, cta, P) : units, u1 = cta < P ? skk_unit_lo(units, cta + 1, P) : units;
  const int acc_cols = MB * SK_BN;                 // TMEM columns of one accumulator buffer

  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB); }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < stages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 4); }
    fence_barrier_init();
  }
  if (warp == 2) { tmem_alloc(tmem_slot, 4 * SK_BN); tmem_relinquish(); }
  pdl_launch_dependents();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

  if (warp == 0 && lane == 0) {
    // frozen weights do not depend on the previous kernel: the first ring of tiles is requested before pdl_wait()
    int n_pre = 0;
    if (p.b_frozen) {
      for (int u = u0; u < u1 && n_pre < stages; ++u, ++n_pre) {
        mbar_expect_tx(&full_bar[n_pre], stage_bytes);
        tma_load_2d(smem + n_pre * stage_bytes + a_bytes, &tmB, (u % nkb) * BK, (u / nkb) * SK_BN, &full_bar[n_pre]);
      }
    }
    pdl_wait();
    uint32_t it = 0;
    for (int u = u0; u < u1; ++u, ++it) {
      const int s = it % stages;
      uint8_t* sa = smem + s * stage_bytes;
      const int k0 = (u % nkb) * BK;
      if ((int)it >= n_pre) {
        mbar_wait(&empty_bar[s], ((it / stages) & 1) ^ 1);
        mbar_expect_tx(&full_bar[s], stage_bytes);
        tma_load_2d(sa + a_bytes, &tmB, k0, (u / nkb) * SK_BN, &full_bar[s]);
      }
      tma_load_2d(sa, &tmA, k0, 0, &full_bar[s]);
      if (MB == 2) tma_load_2d(sa + a_box_bytes, &tmA, k0, BM, &full_bar[s]);
    }
  } else if (warp == 1) {
    pdl_wait();
    const uint32_t idesc = umma_idesc_bf16(BM, SK_BN, 0, 0);
    uint32_t it = 0, seg = 0;
    for (int u = u0; u < u1; ++u, ++it) {
      const int kb = u % nkb;
      const bool first = (u == u0) || kb == 0, last = (u == u1 - 1) || kb == nkb - 1;
      const int buf = seg & 1;
      if (first) {
        mbar_wait(&tempty_bar[buf], ((seg >> 1) & 1) ^ 1);
        tc_fence_after();
      }
      const int s = it % stages;
      mbar_wait(&full_bar[s], (it / stages) & 1);
      tc_fence_after();
      const uint32_t sa = smem_u32(smem + s * stage_bytes);
      const uint32_t sb = sa + a_bytes;
      if (elect_one()) {
        for (int rb = 0; rb < MB; ++rb) {
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)
            tc_mma_ss(tmem_base + buf * acc_cols + rb * SK_BN, umma_desc_sw128(sa + rb * a_box_bytes + k * 32, 16, 1024),
                      umma_desc_sw128(sb + k * 32, 16, 1024), idesc, (!first) || k != 0);
        }
        tc_commit(&empty_bar[s]);
        if (last) tc_commit(&tfull_bar[buf]);
      }
      __syncwarp();
      if (last) ++seg;
    }
  } else if (warp >= 4) {
    pdl_wait();
    const int q = warp & 3, t = threadIdx.x - 128;
    float* stgw = stg + q * 1024;
    const size_t slice = (size_t)MB * BM * SK_BN;           // floats per partial-tile slice
    uint32_t seg = 0;
    int u = u0;
    while (u < u1) {
      const int tile = u / nkb;
      const int end = (tile + 1) * nkb < u1 ? (tile + 1) * nkb : u1;
      const int buf = seg & 1;
      const int first_owner = skk_unit_owner(units, tile * nkb, P), nseg = skk_unit_owner(units, (tile + 1) * nkb - 1, P) - first_owner + 1;
      const int n0 = tile * SK_BN;
      mbar_wait(&tfull_bar[buf], (seg >> 1) & 1);
      tc_fence_after();
      if (nseg == 1) {
        // the whole tile was this CTA's: epilogue straight from TMEM
        for (int cc = 0; cc < MB * (SK_BN / 64); ++cc) {
          const int rb = cc / (SK_BN / 64), c = cc % (SK_BN / 64);
          const int row = rb * BM + q * 32 + lane;
          uint32_t r0[32], r1[32];
          const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + buf * acc_cols + rb * SK_BN + c * 64;
          tmem_ld32(taddr, r0);
          tmem_ld32(taddr + 32, r1);
          tmem_wait_ld();
          float v[64];
#pragma unroll
          for (int j = 0; j < 32; ++j) { v[j] = __uint_as_float(r0[j]); v[j + 32] = __uint_as_float(r1[j]); }
          if (row < p.M) epilogue_chunk<EPI>(p, row, n0 + c * 64, v);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty_bar[buf]);
      } else {
        // ---- park this segment's partial tile: TMEM -> transpose -> rows of 128 contiguous bytes ----
        float* mine = ws + ((size_t)tile * maxseg + (cta - first_owner)) * slice;
        for (int rb = 0; rb < MB; ++rb) {
          const int nrow = p.M - rb * BM - q * 32;              // valid rows of this warp's 32-row block
          for (int g = 0; g < SK_BN / 32; ++g) {
            uint32_t rg[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + buf * acc_cols + rb * SK_BN + g * 32, rg);
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) stgw[lane * 32 + (i ^ lane)] = __uint_as_float(rg[i]);
            __syncwarp();
            float* dst = mine + ((size_t)rb * BM + q * 32) * SK_BN + g * 32 + lane;
#pragma unroll 8
            for (int rr = 0; rr < 32; ++rr)
              if (rr < nrow) __stcg(dst + (size_t)rr * SK_BN, stgw[rr * 32 + (lane ^ rr)]);
            __syncwarp();
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty_bar[buf]);
        // ---- arrival; the CTA that completes the tile finishes it ----
        __threadfence();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (t == 0) {
          const unsigned int old = atomicAdd(counters + tile, 1u);
          const int fin = old == (unsigned int)(nseg - 1);
          if (fin) counters[tile] = 0;                          // every segment has arrived: ready for the next launch
          *last_flag = fin;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        const bool fin = *reinterpret_cast<volatile int*>(last_flag) != 0;
        asm volatile("bar.sync 1, 128;" ::: "memory");            // last_flag is rewritten by the next segment
        if (fin) {
          __threadfence();
          const float* base = ws + (size_t)tile * maxseg * slice;
          for (int cc = 0; cc < MB * (SK_BN / 64); ++cc) {
            const int rb = cc / (SK_BN / 64), c = cc % (SK_BN / 64);
            const int row = rb * BM + q * 32 + lane;
            if (row < p.M) {
              // thread = row: 16 independent 16-byte loads per slice in flight, the slices summed in CTA order (deterministic)
              const float* src = base + (size_t)row * SK_BN + c * 64;
              float v[64];
#pragma unroll
              for (int i = 0; i < 64; ++i) v[i] = 0.f;
              for (int s2 = 0; s2 < nseg; ++s2) {
                float4 x[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) x[i] = __ldcg(reinterpret_cast<const float4*>(src + (size_t)s2 * slice) + i);
#pragma unroll
                for (int i = 0; i < 16; ++i) { v[4 * i] += x[i].x; v[4 * i + 1] += x[i].y; v[4 * i + 2] += x[i].z; v[4 * i + 3] += x[i].w; }
              }
              epilogue_chunk<EPI>(p, row, n0 + c * 64, v);
            }
          }
        }
      }
      ++seg;
      u = end;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 4 * SK_BN);
  }
}

// ------------------------------------------------------------------------------------------------
// Epilogue of the CTA-pair kernel: all global traffic is TMA.  An epilogue warp owns 32 accumulator rows (TMEM
// lanes; thread = row) and works in units of 64 accumulator columns.  Per unit it has an 8 KB shared-memory
// tile pair laid out exactly as the tensor maps' boxes (128-byte rows, 128B swizzle => row-per-lane accesses
// are bank-conflict free): tensors that are read (residual stream, h12 for the SwiGLU backward, the fp32
// output when accumulating) are prefetched one unit ahead by cp.async.bulk.tensor loads, combined with the
// accumulator row in place, and written back by a bulk tensor store.  No thread issues a global ld/st, ragged
// M is clipped by the tensor map, and the tile buffers are double buffered so loads, math and stores overlap.
// ------------------------------------------------------------------------------------------------
constexpr int EPI_BOX_BYTES = 4096;                // one box: 32 rows x 128 B
constexpr int EPI_WARP_BYTES = 2 * EPI_BOX_BYTES;  // two boxes per epilogue warp
constexpr int PAIR_EPI_WARPS = 8;                  // two warps per TMEM lane quadrant
constexpr int PAIR_THREADS = 128 + 32 * PAIR_EPI_WARPS;

__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, uint32_t src_smem, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(m), "r"(src_smem), "r"(c0),
               "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_load_2d_s(uint32_t dst_smem, const CUtensorMap* m, int c0, int c1, uint64_t* bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst_smem),
               "l"(m), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 u;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w) : "r"(addr) : "memory");
  return u;
}
__device__ __forceinline__ void sts128(uint32_t addr, const uint4& u) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(u.x), "r"(u.y), "r"(u.z), "r"(u.w) : "memory");
}

// byte offset of 16-byte chunk c of row r in a box with 128-byte rows (SWIZZLE_128B) / 64-byte rows (SWIZZLE_64B)
__device__ __forceinline__ uint32_t box_off128(int r, int c) { return (uint32_t)(r * 128 + (((c ^ r) & 7) << 4)); }
__device__ __forceinline__ uint32_t box_off64(int r, int c) { return (uint32_t)(r * 64 + (((c ^ (r >> 1)) & 3) << 4)); }

// round two fp32 values to bf16 and back with one packed convert (ALU pipe) instead of two F2F (XU pipe)
__device__ __forceinline__ void bf16_round2(float& a, float& b) {
  const uint32_t u = pack_bf16x2(a, b);
  a = __uint_as_float(u << 16);
  b = __uint_as_float(u & 0xffff0000u);
}

// accumulator columns per epilogue unit: every unit moves 4 KB boxes
template <int EPI> __host__ __device__ constexpr int epi_unit_cols() {
  return (EPI == DH_EPI_F32 || EPI == DH_EPI_RESID_F32 || EPI == DH_EPI_SWIGLU_BWD) ? 32 : 64;
}
template <int EPI> __device__ __forceinline__ bool epi_loads(const GemmParams& p) {
  if constexpr (EPI == DH_EPI_RESID_F32 || EPI == DH_EPI_RESID_BF16 || EPI == DH_EPI_SWIGLU_BWD) return true;
  if constexpr (EPI == DH_EPI_F32) return p.accumulate != 0;
  return false;
}

// lane 0: prefetch the box one unit reads
template <int EPI>
__device__ __forceinline__ void epi_issue_load(const CUtensorMap* tmAux, uint32_t buf, uint64_t* bar, int row0, int col0) {
  mbar_expect_tx(bar, EPI_BOX_BYTES);
  tma_load_2d_s(buf, tmAux, EPI == DH_EPI_SWIGLU_BWD ? 2 * col0 : col0, row0, bar);
}

// lane 0: write the unit's box(es) back
template <int EPI>
__device__ __forceinline__ void epi_issue_store(const GemmParams& p, const CUtensorMap* tmOut, const CUtensorMap* tmOut2, uint32_t buf,
                                                uint32_t buf2, int row0, int col0) {
  if constexpr (EPI == DH_EPI_SWIGLU) {
    if (p.out) tma_store_2d(tmOut, buf, col0, row0);
    tma_store_2d(tmOut2, buf2, col0 >> 1, row0);
  } else if constexpr (EPI == DH_EPI_SWIGLU_BWD) {
    tma_store_2d(tmOut, buf, 2 * col0, row0);
  } else {
    tma_store_2d(tmOut, buf, col0, row0);
  }
  tma_store_commit();
}

template <int NCOL>
__device__ __forceinline__ void box_put_bf16(uint32_t box, int lane, const float (&f)[NCOL]) {
#pragma unroll
  for (int c = 0; c < NCOL / 8; ++c) {
    uint4 u;
    u.x = pack_bf16x2(f[8 * c + 0], f[8 * c + 1]); u.y = pack_bf16x2(f[8 * c + 2], f[8 * c + 3]);
    u.z = pack_bf16x2(f[8 * c + 4], f[8 * c + 5]); u.w = pack_bf16x2(f[8 * c + 6], f[8 * c + 7]);
    sts128(box + box_off128(lane, c), u);
  }
}

struct RopeRegs { uint4 c[4], s[4]; };   // cos/sin[pos][0..32) as packed bf16 (table halves are identical)

// thread = accumulator row `lane` of the unit, v = its epi_unit_cols() columns starting at col0; boxes are updated in place
template <int EPI, int NCOL>
__device__ __forceinline__ void epi_compute(const GemmParams& p, uint32_t buf, uint32_t buf2, int lane, int col0, float (&v)[NCOL],
                                            const RopeRegs& rope) {
  if constexpr (EPI == DH_EPI_BF16) {
#pragma unroll
    for (int j = 0; j < 64; ++j) v[j] *= p.alpha;
    if (p.col_bias) {
#pragma unroll
      for (int j = 0; j < 64; ++j) v[j] = bf16_round(v[j]) + __ldg(p.col_bias + col0 + j);
    }
    if (p.col_scale) {
#pragma unroll
      for (int j = 0; j < 64; ++j) v[j] *= __ldg(p.col_scale + col0 + j);
    }
    box_put_bf16<64>(buf, lane, v);
  } else if constexpr (EPI == DH_EPI_F32) {
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const uint32_t slot = buf + box_off128(lane, c);
      float4 a = make_float4(v[4 * c] * p.alpha, v[4 * c + 1] * p.alpha, v[4 * c + 2] * p.alpha, v[4 * c + 3] * p.alpha);
      if (p.accumulate) {
        const uint4 old = lds128(slot);
        a.x += __uint_as_float(old.x); a.y += __uint_as_float(old.y); a.z += __uint_as_float(old.z); a.w += __uint_as_float(old.w);
      }
      sts128(slot, make_uint4(__float_as_uint(a.x), __float_as_uint(a.y), __float_as_uint(a.z), __float_as_uint(a.w)));
    }
  } else if constexpr (EPI == DH_EPI_RESID_F32) {
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const uint32_t slot = buf + box_off128(lane, c);
      const uint4 x = lds128(slot);
      float a0 = v[4 * c], a1 = v[4 * c + 1], a2 = v[4 * c + 2], a3 = v[4 * c + 3];
      bf16_round2(a0, a1);   // bf16(acc) is what the reference adds to the fp32 stream
      bf16_round2(a2, a3);
      sts128(slot, make_uint4(__float_as_uint(__uint_as_float(x.x) + a0), __float_as_uint(__uint_as_float(x.y) + a1),
                              __float_as_uint(__uint_as_float(x.z) + a2), __float_as_uint(__uint_as_float(x.w) + a3)));
    }
  } else if constexpr (EPI == DH_EPI_RESID_BF16) {
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const uint32_t slot = buf + box_off128(lane, c);
      float x[8];
      unpack_bf16x8(lds128(slot), x);
#pragma unroll
      for (int k = 0; k < 8; k += 2) {
        float a0 = v[8 * c + k], a1 = v[8 * c + k + 1];
        bf16_round2(a0, a1);
        x[k] += a0; x[k + 1] += a1;
      }
      sts128(slot, pack_bf16x8(x));
    }
  } else if constexpr (EPI == DH_EPI_QKV_ROPE) {
    const int slot = (col0 >> 6) % (p.q_per_kv + 2);
    if (slot <= p.q_per_kv) {   // q heads and the k head rotate; v passes through
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t cw[4] = {rope.c[i].x, rope.c[i].y, rope.c[i].z, rope.c[i].w};
        const uint32_t sw[4] = {rope.s[i].x, rope.s[i].y, rope.s[i].z, rope.s[i].w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float2 cc = unpack_bf16x2(cw[k]), ss = unpack_bf16x2(sw[k]);
          const int j = 8 * i + 2 * k;
          float a0 = v[j], a1 = v[j + 1], b0 = v[j + 32], b1 = v[j + 33];
          bf16_round2(a0, a1);
          bf16_round2(b0, b1);
          v[j] = a0 * cc.x - b0 * ss.x;      v[j + 32] = b0 * cc.x + a0 * ss.x;
          v[j + 1] = a1 * cc.y - b1 * ss.y;  v[j + 33] = b1 * cc.y + a1 * ss.y;
        }
      }
    }
    box_put_bf16<64>(buf, lane, v);
  } else if constexpr (EPI == DH_EPI_SWIGLU) {
    // raw h12 (64 cols) in the first box, act = silu(h1) * h2 (32 cols, 64-byte rows, SWIZZLE_64B) in the second
    if (p.out) box_put_bf16<64>(buf, lane, v);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float a[8];
#pragma unroll
      for (int k = 0; k < 8; k += 2) {
        float h1a = v[8 * c + k], h1b = v[8 * c + k + 1], h2a = v[8 * c + k + 32], h2b = v[8 * c + k + 33];
        bf16_round2(h1a, h1b);
        bf16_round2(h2a, h2b);
        a[k] = h1a * fast_sigmoid(h1a) * h2a;
        a[k + 1] = h1b * fast_sigmoid(h1b) * h2b;
      }
      sts128(buf2 + box_off64(lane, c), pack_bf16x8(a));
    }
  } else if constexpr (EPI == DH_EPI_SWIGLU_BWD) {
    // acc = d(act) columns col0..col0+31; the box holds h12 columns 2*col0 .. 2*col0+63 = [h1 x32 | h2 x32]
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const uint32_t s1 = buf + box_off128(lane, c), s2 = buf + box_off128(lane, c + 4);
      float h1[8], h2[8], d1[8], d2[8];
      unpack_bf16x8(lds128(s1), h1);
      unpack_bf16x8(lds128(s2), h2);
#pragma unroll
      for (int k = 0; k < 8; k += 2) {
        float da0 = v[8 * c + k], da1 = v[8 * c + k + 1];
        bf16_round2(da0, da1);
        const float sg0 = fast_sigmoid(h1[k]), sg1 = fast_sigmoid(h1[k + 1]);
        d1[k] = da0 * h2[k] * (sg0 * (1.f + h1[k] * (1.f - sg0)));
        d2[k] = da0 * (h1[k] * sg0);
        d1[k + 1] = da1 * h2[k + 1] * (sg1 * (1.f + h1[k + 1] * (1.f - sg1)));
        d2[k + 1] = da1 * (h1[k + 1] * sg1);
      }
      sts128(s1, pack_bf16x8(d1));
      sts128(s2, pack_bf16x8(d2));
    }
  }
}

// ------------------------------------------------------------------------------------------------
// CTA-pair kernel (cluster of 2, tcgen05 cta_group::2): one 256 x 256 output tile per pair.  CTA r of the
// pair stages A rows [128 r, 128 r + 128) and B rows (= output columns) [128 r, 128 r + 128) of the tile
// and owns accumulator rows [128 r, +128) x 256 columns in its own TMEM; the leader's MMA thread drives
// both tensor cores, so every B byte is fetched into shared memory once per pair instead of once per SM.
// K-major operands only (dgrads use pre-transposed frozen weights).
// Roles (384 threads): warp 0 = TMA producer, warp 1 = MMA issuer, warp 2 = TMEM allocator, warps 4-11 = epilogue
// (warp w owns TMEM lanes 32*(w%4).., and the units of a tile alternate between the two warps of a quadrant).
// ------------------------------------------------------------------------------------------------
struct PairCfg {
  static constexpr int BN = 256;
  static constexpr int A_BYTES = 128 * BK * 2;
  static constexpr int B_BYTES = 128 * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = 5;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + PAIR_EPI_WARPS * EPI_WARP_BYTES + 1024 + 512;
};

struct PairMaps {
  CUtensorMap A, B, A2, B2, out, out2, aux;
};

template <int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(PAIR_THREADS, 1)
gemm_pair_kernel(const __grid_constant__ PairMaps tm, const GemmParams p) {
  using Cfg = PairCfg;
  constexpr int STAGES = Cfg::STAGES;
  constexpr int BN = Cfg::BN;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* epi_base = smem + STAGES * Cfg::STAGE_BYTES;   // 8 epilogue warps x 2 x 4 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_base + PAIR_EPI_WARPS * EPI_WARP_BYTES);
  uint64_t* full_bar = bars;                     // [STAGES]  TMA (both CTAs) -> MMA       (leader's copy is used)
  uint64_t* empty_bar = bars + STAGES;           // [STAGES]  MMA -> TMA                   (multicast to both)
  uint64_t* tfull_bar = bars + 2 * STAGES;       // [2]       MMA -> epilogue              (multicast to both)
  uint64_t* tempty_bar = bars + 2 * STAGES + 2;  // [2]       epilogue (both CTAs) -> MMA  (leader's copy is used)
  uint64_t* eload_bar = bars + 2 * STAGES + 4;   // [8][2]    epilogue box prefetch, per warp and buffer
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4 + 2 * PAIR_EPI_WARPS);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // shfl: a warp-uniform value for the compiler
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int pair = blockIdx.x >> 1;
  const int npairs = gridDim.x >> 1;
  const int nkb1 = (p.K1 + BK - 1) / BK;
  const int nkb = nkb1 + (p.K2 + BK - 1) / BK;
  pdl_launch_dependents();

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm.A);
    tma_prefetch_desc(&tm.B);
    if (p.K2 > 0) { tma_prefetch_desc(&tm.A2); tma_prefetch_desc(&tm.B2); }
  }
  if (warp == 3 && lane == 0) {
    if (p.out || EPI != DH_EPI_SWIGLU) tma_prefetch_desc(&tm.out);
    if constexpr (EPI == DH_EPI_SWIGLU) tma_prefetch_desc(&tm.out2);
    if (epi_loads<EPI>(p)) tma_prefetch_desc(&tm.aux);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 2 * PAIR_EPI_WARPS); }
    for (int i = 0; i < 2 * PAIR_EPI_WARPS; ++i) mbar_init(&eload_bar[i], 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc2(tmem_slot, 512);
    tmem_relinquish2();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  pdl_wait();          // the prologue above overlapped the previous kernel's tail; global memory is read from here on
  int tiles_m = p.tiles_m;
  if (p.m_limit) { const int t = (__ldg(p.m_limit) + 255) / 256; tiles_m = t < tiles_m ? t : tiles_m; }
  const int num_tiles = tiles_m * p.tiles_n;

  if (warp == 0 && lane == 0) {
    // ===================== TMA producer (each CTA loads its own halves) =====================
    uint32_t it = 0;
    for (int tile = pair; tile < num_tiles; tile += npairs) {
      int tm_, tn_;
      pair_tile_coords(tile, tiles_m, p.tiles_n, p.band_m, tm_, tn_);
      const int m0 = tm_ * 256 + (int)rank * 128;
      const int n0 = tn_ * BN + (int)rank * 128;
      for (int kb = 0; kb < nkb; ++kb, ++it) {
        const int s = it % STAGES;
        mbar_wait(&empty_bar[s], ((it / STAGES) & 1) ^ 1);
        uint8_t* sa = smem + s * Cfg::STAGE_BYTES;
        uint8_t* sb = sa + Cfg::A_BYTES;
        if (rank == 0) mbar_expect_tx(&full_bar[s], 2 * Cfg::STAGE_BYTES);
        const uint32_t bar = mapa_shared(smem_u32(&full_bar[s]), 0);
        const bool second = kb >= nkb1;
        const int k0 = (second ? kb - nkb1 : kb) * BK;
        tma_load_2d_pair(sa, second ? &tm.A2 : &tm.A, k0, m0, bar);
        tma_load_2d_pair(sb, second ? &tm.B2 : &tm.B, k0, n0, bar);
      }
    }
  } else if (warp == 1 && rank == 0) {
    // ===================== MMA issuer (leader CTA only) =====================
    constexpr uint32_t idesc = umma_idesc_bf16(256, BN, 0, 0);
    uint32_t it = 0, tcount = 0;
    for (int tile = pair; tile < num_tiles; tile += npairs, ++tcount) {
      const int acc = tcount & 1;
      mbar_wait(&tempty_bar[acc], ((tcount >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * BN;
      for (int kb = 0; kb < nkb; ++kb, ++it) {
        const int s = it % STAGES;
        mbar_wait(&full_bar[s], (it / STAGES) & 1);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + s * Cfg::STAGE_BYTES);
        const uint32_t sb = sa + Cfg::A_BYTES;
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)
            tc_mma_ss_pair(d_tmem, umma_desc_sw128(sa + k * 32, 16, 1024), umma_desc_sw128(sb + k * 32, 16, 1024), idesc, (kb | k) != 0);
          tc_commit_pair(&empty_bar[s], 3);
          if (kb == nkb - 1) tc_commit_pair(&tfull_bar[acc], 3);
        }
        __syncwarp();
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue (each CTA drains its own 128 accumulator rows) =====================
    constexpr int UC = epi_unit_cols<EPI>();     // accumulator columns per unit
    constexpr int UPW = BN / UC / 2;             // units per warp per tile
    constexpr bool SINGLE_BUF = (EPI == DH_EPI_SWIGLU);   // uses both boxes for one unit
    const int q = warp & 3, half = (warp - 4) >> 2;
    const uint32_t ebuf = smem_u32(epi_base + (warp - 4) * EPI_WARP_BYTES);
    uint64_t* lbar = eload_bar + 2 * (warp - 4);
    const bool loads = epi_loads<EPI>(p);
    const int my_tiles = pair < num_tiles ? (num_tiles - pair + npairs - 1) / npairs : 0;
    const int n_units = my_tiles * UPW;
    // unit u of this warp = column slice 2 * (u % UPW) + half of this pair's (u / UPW)-th tile
    auto unit_row0 = [&](int u) {
      int tm_, tn_;
      pair_tile_coords(pair + (u / UPW) * npairs, tiles_m, p.tiles_n, p.band_m, tm_, tn_);
      return tm_ * 256 + (int)rank * 128 + q * 32;
    };
    auto unit_col0 = [&](int u) {
      int tm_, tn_;
      pair_tile_coords(pair + (u / UPW) * npairs, tiles_m, p.tiles_n, p.band_m, tm_, tn_);
      return tn_ * BN + (2 * (u % UPW) + half) * UC;
    };
    if (loads && lane == 0 && n_units > 0) epi_issue_load<EPI>(&tm.aux, ebuf, &lbar[0], unit_row0(0), unit_col0(0));
    RopeRegs rope;
    for (int u = 0; u < n_units; ++u) {
      const int tcount = u / UPW, acc = tcount & 1, c = 2 * (u % UPW) + half;
      const int row0 = unit_row0(u), col0 = unit_col0(u);
      const uint32_t buf = SINGLE_BUF ? ebuf : ebuf + (u & 1) * EPI_BOX_BYTES;
      if (u % UPW == 0) {
        mbar_wait(&tfull_bar[acc], (tcount >> 1) & 1);
        tc_fence_after();
        if constexpr (EPI == DH_EPI_QKV_ROPE) {
          const int row = row0 + lane < p.M ? row0 + lane : p.M - 1;
          const int pos = p.positions ? __ldg(p.positions + row) : (row % p.seq_len);
          const uint4* c4 = reinterpret_cast<const uint4*>(p.rope_cos + (long long)pos * 64);
          const uint4* s4 = reinterpret_cast<const uint4*>(p.rope_sin + (long long)pos * 64);
#pragma unroll
          for (int i = 0; i < 4; ++i) { rope.c[i] = __ldg(c4 + i); rope.s[i] = __ldg(s4 + i); }
        }
      }
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN + c * UC;
      float v[UC];
      {
        uint32_t r0[32];
        tmem_ld32(taddr, r0);
        if constexpr (UC == 64) {
          uint32_t r1[32];
          tmem_ld32(taddr + 32, r1);
          tmem_wait_ld();
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j + 32 * (UC == 64)] = __uint_as_float(r1[j]);
        } else {
          tmem_wait_ld();
        }
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r0[j]);
      }
      if (lane == 0) {
        if (loads) {
          tma_store_wait_read<0>();   // the other box's store (unit u-1) has left shared memory: refill it
          if (u + 1 < n_units)
            epi_issue_load<EPI>(&tm.aux, ebuf + ((u + 1) & 1) * EPI_BOX_BYTES, &lbar[(u + 1) & 1], unit_row0(u + 1), unit_col0(u + 1));
        } else if (SINGLE_BUF) {
          tma_store_wait_read<0>();
        } else {
          tma_store_wait_read<1>();   // this box's previous store (unit u-2) has left shared memory
        }
      }
      if (u % UPW == UPW - 1) tc_fence_before();   // this warp has read its share of the accumulator: hand it back
      __syncwarp();
      if (u % UPW == UPW - 1 && lane == 0) mbar_arrive_cluster(mapa_shared(smem_u32(&tempty_bar[acc]), 0));
      if (loads) mbar_wait(&lbar[u & 1], (u >> 1) & 1);
      epi_compute<EPI, UC>(p, buf, ebuf + EPI_BOX_BYTES, lane, col0, v, rope);
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) epi_issue_store<EPI>(p, &tm.out, &tm.out2, buf, ebuf + EPI_BOX_BYTES, row0, col0);
    }
    if (lane == 0) tma_store_wait_all();
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc2(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------------
// host
// ------------------------------------------------------------------------------------------------
static int make_operand_map(CUtensorMap* map, const void* base, long long ld, int mn_extent, int k_extent, int mn_major,
                            int box_mn) {
  if (!mn_major) {  // [mn, k] row-major: inner dim k
    uint64_t dims[2] = {(uint64_t)k_extent, (uint64_t)mn_extent};
    uint64_t strides[1] = {(uint64_t)ld * 2};
    uint32_t box[2] = {64, (uint32_t)box_mn};
    return make_tensor_map(map, base, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
  }
  uint64_t dims[2] = {(uint64_t)mn_extent, (uint64_t)k_extent};  // [k, mn] row-major: inner dim mn
  uint64_t strides[1] = {(uint64_t)ld * 2};
  uint32_t box[2] = {64, 64};
  return make_tensor_map(map, base, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

static GemmParams make_params(const dh_gemm_args* a) {
  GemmParams p;
  p.M = a->M; p.N = a->N; p.K1 = a->K1; p.K2 = a->K2;
  p.tiles_m = p.tiles_n = 0;
  p.a_mn = a->a_layout; p.b_mn = a->b_layout; p.a2_mn = a->a2_layout; p.b2_mn = a->b2_layout;
  p.accumulate = a->accumulate;
  p.alpha = a->alpha;
  p.seq_len = a->seq_len > 0 ? a->seq_len : 1;
  p.q_per_kv = a->q_per_kv;
  p.out = a->out; p.ldo = a->ldo;
  p.out2 = a->out2; p.ldo2 = a->ldo2;
  p.aux = a->aux; p.ldaux = a->ldaux;
  p.col_scale = a->col_scale; p.col_bias = a->col_bias;
  p.rope_cos = reinterpret_cast<const bf16*>(a->rope_cos);
  p.rope_sin = reinterpret_cast<const bf16*>(a->rope_sin);
  p.positions = a->positions;
  p.m_limit = a->m_limit;
  p.splits = 1; p.split_stride = 0;
  p.sk_stages = 0; p.sk_partial = 0; p.b_frozen = 0;
  return p;
}

template <int BN, int EPI>
static int launch_gemm(const dh_gemm_args* a, cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  static bool attr_set = false;
  if (!attr_set) {
    DH_CUDA(cudaFuncSetAttribute(gemm_kernel<BN, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    attr_set = true;
  }
  CUtensorMap tmA, tmB, tmA2, tmB2;
  if (make_operand_map(&tmA, a->A, a->lda, a->M, a->K1, a->a_layout, BM)) return -3;
  if (make_operand_map(&tmB, a->B, a->ldb, a->N, a->K1, a->b_layout, BN)) return -3;
  if (a->K2 > 0) {
    if (make_operand_map(&tmA2, a->A2, a->lda2, a->M, a->K2, a->a2_layout, BM)) return -3;
    if (make_operand_map(&tmB2, a->B2, a->ldb2, a->N, a->K2, a->b2_layout, BN)) return -3;
  } else {
    tmA2 = tmA;
    tmB2 = tmB;
  }
  GemmParams p = make_params(a);
  p.tiles_m = (a->M + BM - 1) / BM;
  p.tiles_n = a->N / BN;
  if (a->split_k > 1) {
    const int nkb = (a->K1 + BK - 1) / BK + (a->K2 + BK - 1) / BK;
    p.splits = a->split_k < nkb ? a->split_k : nkb;
    p.split_stride = (long long)a->M * a->ldo;
  }
  const int tiles = p.tiles_m * p.tiles_n * p.splits;
  const int grid = tiles < sm_count() ? tiles : sm_count();
  DH_CUDA(launch_pdl(gemm_kernel<BN, EPI>, dim3(grid), dim3(GEMM_THREADS), Cfg::SMEM_BYTES, stream, tmA, tmB, tmA2, tmB2, p));
  return 0;
}

// [rows, cols] row-major tensor of `esize`-byte elements, box = {box_cols, 32 rows}
static int make_epi_map(CUtensorMap* map, const void* base, long long ld, long long rows, long long cols, int esize, int box_cols) {
  uint64_t dims[2] = {(uint64_t)cols, (uint64_t)rows};
  uint64_t strides[1] = {(uint64_t)ld * esize};
  uint32_t box[2] = {(uint32_t)box_cols, 32};
  const int row_bytes = box_cols * esize;
  return make_tensor_map(map, base, esize == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, dims, strides, box,
                         row_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B);
}

// can the epilogue tensors of this call be described by tensor maps (16-byte aligned bases and row pitches)?
static bool pair_epilogue_ok(const dh_gemm_args* a) {
  auto ok = [](const void* ptr, long long ld, int esize) { return ((uintptr_t)ptr & 15) == 0 && (ld * esize) % 16 == 0; };
  switch (a->epilogue) {
    case DH_EPI_BF16: case DH_EPI_QKV_ROPE: return ok(a->out, a->ldo, 2);
    case DH_EPI_F32: return ok(a->out, a->ldo, 4);
    case DH_EPI_RESID_F32: return ok(a->out, a->ldo, 4) && ok(a->aux, a->ldaux, 4);
    case DH_EPI_RESID_BF16: return ok(a->out, a->ldo, 2) && ok(a->aux, a->ldaux, 2);
    case DH_EPI_SWIGLU: return (!a->out || ok(a->out, a->ldo, 2)) && ok(a->out2, a->ldo2, 2);
    case DH_EPI_SWIGLU_BWD: return ok(a->out, a->ldo, 2) && ok(a->aux, a->ldaux, 2);
    default: return false;
  }
}

template <int EPI>
static int launch_gemm_pair(const dh_gemm_args* a, cudaStream_t stream) {
  using Cfg = PairCfg;
  static bool attr_set = false;
  if (!attr_set) {
    DH_CUDA(cudaFuncSetAttribute(gemm_pair_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    attr_set = true;
  }
  PairMaps tm;
  if (make_operand_map(&tm.A, a->A, a->lda, a->M, a->K1, 0, 128)) return -3;
  if (make_operand_map(&tm.B, a->B, a->ldb, a->N, a->K1, 0, 128)) return -3;
  if (a->K2 > 0) {
    if (make_operand_map(&tm.A2, a->A2, a->lda2, a->M, a->K2, 0, 128)) return -3;
    if (make_operand_map(&tm.B2, a->B2, a->ldb2, a->N, a->K2, 0, 128)) return -3;
  } else {
    tm.A2 = tm.A;
    tm.B2 = tm.B;
  }
  const long long M = a->M, N = a->N;
  tm.out = tm.out2 = tm.aux = tm.A;   // placeholders for the maps this epilogue never touches
  if constexpr (EPI == DH_EPI_BF16 || EPI == DH_EPI_QKV_ROPE) {
    if (make_epi_map(&tm.out, a->out, a->ldo, M, N, 2, 64)) return -3;
  } else if constexpr (EPI == DH_EPI_F32) {
    if (make_epi_map(&tm.out, a->out, a->ldo, M, N, 4, 32)) return -3;
    tm.aux = tm.out;   // accumulate reads what it overwrites
  } else if constexpr (EPI == DH_EPI_RESID_F32) {
    if (make_epi_map(&tm.out, a->out, a->ldo, M, N, 4, 32)) return -3;
    if (make_epi_map(&tm.aux, a->aux, a->ldaux, M, N, 4, 32)) return -3;
  } else if constexpr (EPI == DH_EPI_RESID_BF16) {
    if (make_epi_map(&tm.out, a->out, a->ldo, M, N, 2, 64)) return -3;
    if (make_epi_map(&tm.aux, a->aux, a->ldaux, M, N, 2, 64)) return -3;
  } else if constexpr (EPI == DH_EPI_SWIGLU) {
    if (a->out && make_epi_map(&tm.out, a->out, a->ldo, M, N, 2, 64)) return -3;
    if (make_epi_map(&tm.out2, a->out2, a->ldo2, M, N / 2, 2, 32)) return -3;
  } else if constexpr (EPI == DH_EPI_SWIGLU_BWD) {
    if (make_epi_map(&tm.out, a->out, a->ldo, M, 2 * N, 2, 64)) return -3;
    if (make_epi_map(&tm.aux, a->aux, a->ldaux, M, 2 * N, 2, 64)) return -3;
  }
  GemmParams p = make_params(a);
  p.tiles_m = (a->M + 255) / 256;
  p.tiles_n = a->N / Cfg::BN;
  {
    // band height: the band's A rows (both sources) within ~20 MB of L2, at least one wave of pairs wide when N is narrow
    static long long budget = -1;
    if (budget < 0) { const char* e = getenv("DH_GEMM_BAND_MB"); budget = (e ? atoll(e) : 20) << 20; }   // measured: 16-28 MB best (prefill 246 -> 196 ms)
    const long long a_tile_bytes = 256LL * (a->K1 + a->K2) * 2;
    long long band = budget / a_tile_bytes;
    const long long wave_m = (sm_count() / 2 + p.tiles_n - 1) / p.tiles_n;      // M-tiles one wave of pairs spans when it covers all N-tiles
    if (band < wave_m) band = wave_m;
    if (band < 1) band = 1;
    p.band_m = band >= p.tiles_m ? p.tiles_m : (int)band;
    if (a->m_limit) p.band_m = p.tiles_m;       // the device-side row bound changes tiles_m: keep the plain order there
  }
  const int tiles = p.tiles_m * p.tiles_n;
  const int pairs = sm_count() / 2;
  const int grid = 2 * (tiles < pairs ? tiles : pairs);
  DH_CUDA(launch_pdl(gemm_pair_kernel<EPI>, dim3(grid), dim3(PAIR_THREADS), Cfg::SMEM_BYTES, stream, tm, p));
  return 0;
}

static thread_local int t_b_frozen = 0;   // DH_B_FROZEN of the call being dispatched (dh_gemm strips it from b_layout)

template <int EPI>
static int launch_gemm_skinny(const dh_gemm_args* a, cudaStream_t stream) {
  static bool attr_set = false;
  if (!attr_set) {
    DH_CUDA(cudaFuncSetAttribute(gemm_skinny_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, SK_RING_BYTES + 1024 + 256));
    attr_set = true;
  }
  const int box_m = a->M <= 64 ? 64 : BM;
  CUtensorMap tmA, tmB, tmA2, tmB2;
  if (make_operand_map(&tmA, a->A, a->lda, a->M, a->K1, 0, box_m)) return -3;
  if (make_operand_map(&tmB, a->B, a->ldb, a->N, a->K1, 0, SK_BN)) return -3;
  if (a->K2 > 0) {
    if (make_operand_map(&tmA2, a->A2, a->lda2, a->M, a->K2, 0, box_m)) return -3;
    if (make_operand_map(&tmB2, a->B2, a->ldb2, a->N, a->K2, 0, SK_BN)) return -3;
  } else {
    tmA2 = tmA;
    tmB2 = tmB;
  }
  GemmParams p = make_params(a);
  p.b_frozen = t_b_frozen;
  p.tiles_m = 1;
  p.tiles_n = a->N / SK_BN;
  const int tiles = p.tiles_n;
  const int nkb = (a->K1 + BK - 1) / BK + (a->K2 + BK - 1) / BK;
  // cluster size: a power of two <= 4 that keeps every CTA resident at once (one wave) with >= 2 k-blocks each.
  // Measured on B200 (scripts/skinny_probe.py): a CTA streams one k-block per ~0.4 us whatever the ring depth (the
  // SM's cap on bytes in flight), a cluster adds ~4 us, and 16 clusters of EIGHT 192 KB CTAs do not all fit at once
  // (two per GPC would need 16 free SMs in every GPC), which doubles the time -- hence the cap of 4.
  int S = 1;
  while (S < 4 && tiles * (2 * S) <= sm_count() && nkb / (2 * S) >= 2) S *= 2;
  const int stage_bytes = box_m * BK * 2 * (a->M > BM ? 2 : 1) + SK_B_BYTES;      // two activation boxes per stage for 128 < M <= 256
  int stages = SK_RING_BYTES / stage_bytes;
  if (stages > SK_MAX_STAGES) stages = SK_MAX_STAGES;
  const int per_cta = (nkb + S - 1) / S;
  if (stages > per_cta) stages = per_cta;
  if (a->split_k > 1) {       // partial products for a consumer that sums them: no cluster, any split count with >= 1 k-block each
    S = a->split_k < nkb ? a->split_k : nkb;
    p.sk_partial = 1;
    stages = SK_RING_BYTES / stage_bytes;
    if (stages > SK_MAX_STAGES) stages = SK_MAX_STAGES;
    const int pc = (nkb + S - 1) / S;
    if (stages > pc) stages = pc;
  }
  static int env_s = -2, env_stages = -2;      // developer knobs for A/B runs: DH_SKINNY_S, DH_SKINNY_STAGES
  if (env_s == -2) { const char* e = getenv("DH_SKINNY_S"); env_s = e ? atoi(e) : -1; }
  if (env_stages == -2) { const char* e = getenv("DH_SKINNY_STAGES"); env_stages = e ? atoi(e) : -1; }
  if (env_s > 0 && nkb / env_s >= 1 && !p.sk_partial) S = env_s;
  if (env_stages > 0 && env_stages <= SK_MAX_STAGES && env_stages * stage_bytes <= SK_RING_BYTES) stages = env_stages;
  // the fp32 partial tile and this CTA's reduced units (S > 1) re-use the ring
  const size_t park = (size_t)a->M * SK_PAD * sizeof(float) + (size_t)((a->M * 2 + S - 1) / S) * 64 * sizeof(float);
  while (S > 1 && !p.sk_partial && (size_t)stages * stage_bytes < park) ++stages;
  p.sk_stages = stages;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(S, tiles);
  cfg.blockDim = dim3(GEMM_THREADS);
  cfg.dynamicSmemBytes = (size_t)stages * stage_bytes + 1024 + 256;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = p.sk_partial ? 1 : S;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (pdl_level() >= 2) {
    cfg.numAttrs = 2;
  } else if (pdl_level() == 1) {
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(stream, &st) == cudaSuccess && st == cudaStreamCaptureStatusNone) cfg.numAttrs = 2;
  }
  DH_CUDA(cudaLaunchKernelEx(&cfg, gemm_skinny_kernel<EPI>, tmA, tmB, tmA2, tmB2, p));
  return 0;
}

// workspace of the stream-K decode GEMM: [arrival counters, one per tile, padded to 256 B][tiles][maxseg][row blocks * 128][128] fp32
static size_t streamk_layout(int M, int N, int K, int* maxseg_out) {
  const int tiles = N / SK_BN, nkb = K / BK, units = tiles * nkb;
  int P = sm_count() > 0 ? sm_count() : 148;
  if (units < P) P = units;
  int maxseg = 1;
  for (int t = 0; t < tiles; ++t) {
    const int c = skk_unit_owner(units, (t + 1) * nkb - 1, P) - skk_unit_owner(units, t * nkb, P) + 1;
    if (c > maxseg) maxseg = c;
  }
  if (maxseg_out) *maxseg_out = maxseg;
  const size_t counters = ((size_t)tiles * sizeof(unsigned int) + 255) & ~(size_t)255;
  return counters + (size_t)tiles * maxseg * (M > BM ? 2 : 1) * BM * SK_BN * sizeof(float);
}

template <int EPI>
static int launch_gemm_streamk(const dh_gemm_args* a, cudaStream_t stream) {
  const int box_m = a->M <= 64 ? 64 : BM;
  const int stage_bytes = box_m * BK * 2 * (a->M > BM ? 2 : 1) + SK_B_BYTES;
  int stages = SK_RING_BYTES / stage_bytes;
  if (stages > SK_MAX_STAGES) stages = SK_MAX_STAGES;
  const size_t smem = (size_t)stages * stage_bytes + SKK_STG_BYTES + 1024 + 256;
  static bool attr_set = false;
  if (!attr_set) {
    DH_CUDA(cudaFuncSetAttribute(gemm_streamk_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set = true;
  }
  DH_CHECK(smem <= 227 * 1024, "dh_gemm (stream-K): shared memory %zu", smem);
  int maxseg = 1;
  const size_t need = streamk_layout(a->M, a->N, a->K1, &maxseg);
  DH_CHECK(a->workspace_bytes >= need, "dh_gemm (stream-K): workspace of %zu bytes given, %zu needed (dh_gemm_streamk_workspace_bytes)",
           (size_t)a->workspace_bytes, need);
  CUtensorMap tmA, tmB;
  if (make_operand_map(&tmA, a->A, a->lda, a->M, a->K1, 0, box_m)) return -3;
  if (make_operand_map(&tmB, a->B, a->ldb, a->N, a->K1, 0, SK_BN)) return -3;
  GemmParams p = make_params(a);
  p.b_frozen = t_b_frozen;
  p.tiles_m = 1;
  p.tiles_n = a->N / SK_BN;
  p.sk_stages = stages;
  const int tiles = p.tiles_n;
  unsigned int* counters = reinterpret_cast<unsigned int*>(a->workspace);
  float* ws = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(a->workspace) + (((size_t)tiles * sizeof(unsigned int) + 255) & ~(size_t)255));
  DH_CUDA(launch_pdl(gemm_streamk_kernel<EPI>, dim3(sm_count()), dim3(GEMM_THREADS), smem, stream, tmA, tmB, p, ws, counters, maxseg));
  return 0;
}

static int g_force_single = -1;   // DH_GEMM_SINGLE=1 forces the 1-CTA kernel (A/B testing)

template <int EPI>
static int dispatch_bn(const dh_gemm_args* a, cudaStream_t stream) {
  if (g_force_single < 0) {
    const char* e = getenv("DH_GEMM_SINGLE");
    g_force_single = (e && e[0] == '1') ? 1 : 0;
  }
  // CTA pairs for the big K-major GEMMs of the decoder; the 1-CTA kernel covers small M, MN-major operands, odd N.
  if (!g_force_single && a->a_layout == 0 && a->b_layout == 0 && (a->K2 == 0 || (a->a2_layout == 0 && a->b2_layout == 0)) &&
      a->N % 256 == 0 && a->M >= 1024 && a->split_k <= 1 && pair_epilogue_ok(a))
    return launch_gemm_pair<EPI>(a, stream);
  // decode GEMM with a tile-local epilogue and too few tiles for the machine: stream-K over all SMs (the caller lends a workspace)
  if constexpr (EPI == DH_EPI_SWIGLU) {
    if (a->workspace && a->M <= 2 * BM && a->N % SK_BN == 0 && a->K1 % BK == 0 && a->K2 == 0 && a->a_layout == 0 && a->b_layout == 0 &&
        !a->m_limit && a->N / SK_BN < sm_count())
      return launch_gemm_streamk<EPI>(a, stream);
  }
  // one token per sequence (decode): weight streaming with K split over a cluster
  if (a->M <= 2 * BM && a->N % SK_BN == 0 && a->a_layout == 0 && a->b_layout == 0 && (a->K2 == 0 || (a->a2_layout == 0 && a->b2_layout == 0)) &&
      !a->m_limit && EPI != DH_EPI_SWIGLU_BWD && (EPI != DH_EPI_F32 || !a->accumulate))
    return launch_gemm_skinny<EPI>(a, stream);
  // Tile width: 256 when it divides N and leaves enough tiles to fill the machine, else 128, else 64.
  const long long tiles_m = (a->M + BM - 1) / BM;
  const long long sk = a->split_k > 1 ? a->split_k : 1;
  if (a->N % 256 == 0 && tiles_m * (a->N / 256) * sk >= sm_count()) return launch_gemm<256, EPI>(a, stream);
  if (a->N % 128 == 0 && tiles_m * (a->N / 128) * sk >= sm_count() / 2) return launch_gemm<128, EPI>(a, stream);
  return launch_gemm<64, EPI>(a, stream);
}

}  // namespace dh

extern "C" size_t dh_gemm_streamk_workspace_bytes(int M, int N, int K) {
  if (M <= 0 || N <= 0 || K <= 0 || N % dh::SK_BN || K % dh::BK) return 0;
  return dh::streamk_layout(M, N, K, nullptr);
}

extern "C" int dh_gemm(const dh_gemm_args* args, void* stream_) {
  using namespace dh;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DH_CHECK(args != nullptr, "dh_gemm: null args");
  dh_gemm_args local = *args;
  t_b_frozen = (local.b_layout & DH_B_FROZEN) ? 1 : 0;
  local.b_layout &= ~DH_B_FROZEN;
  const dh_gemm_args* a = &local;
  DH_CHECK(a->M > 0 && a->N > 0 && a->K1 > 0, "dh_gemm: bad shape M=%d N=%d K1=%d", a->M, a->N, a->K1);
  // K is the row pitch of a K-major operand (16-byte aligned rows => K % 8); for MN-major operands (the wgrad forms, K = the token
  // count of a packed batch) K only counts rows and may be anything: TMA zero-fills the rows of the last k-block
  const bool k1_pitch = a->a_layout == 0 || a->b_layout == 0;
  const bool k2_pitch = a->K2 > 0 && (a->a2_layout == 0 || a->b2_layout == 0);
  DH_CHECK(a->N % 64 == 0 && (!k1_pitch || a->K1 % 8 == 0) && (!k2_pitch || a->K2 % 8 == 0),
           "dh_gemm: N %% 64 must be 0, and K %% 8 for K-major operands (N=%d K1=%d K2=%d)", a->N, a->K1, a->K2);
  DH_CHECK(a->A && a->B && (a->out || a->epilogue == DH_EPI_SWIGLU), "dh_gemm: null operand");
  DH_CHECK((a->lda % 8) == 0 && (a->ldb % 8) == 0, "dh_gemm: leading dimensions must be multiples of 8 elements");
  DH_CHECK(((uintptr_t)a->A & 15) == 0 && ((uintptr_t)a->B & 15) == 0, "dh_gemm: operands must be 16-byte aligned");
  if (a->K2 > 0) DH_CHECK(a->A2 && a->B2 && (a->lda2 % 8) == 0 && (a->ldb2 % 8) == 0, "dh_gemm: bad second source");
  if (a->split_k > 1)
    DH_CHECK(a->epilogue == DH_EPI_F32 && !a->accumulate && !a->m_limit, "dh_gemm: split_k needs the plain DH_EPI_F32 epilogue without m_limit");
  switch (a->epilogue) {
    case DH_EPI_BF16: return dispatch_bn<DH_EPI_BF16>(a, stream);
    case DH_EPI_F32: return dispatch_bn<DH_EPI_F32>(a, stream);
    case DH_EPI_RESID_F32:
      DH_CHECK(a->aux, "dh_gemm: residual epilogue needs aux");
      return dispatch_bn<DH_EPI_RESID_F32>(a, stream);
    case DH_EPI_RESID_BF16:
      DH_CHECK(a->aux, "dh_gemm: residual epilogue needs aux");
      return dispatch_bn<DH_EPI_RESID_BF16>(a, stream);
    case DH_EPI_QKV_ROPE:
      DH_CHECK(a->rope_cos && a->rope_sin && a->q_per_kv > 0, "dh_gemm: rope epilogue needs tables and q_per_kv");
      DH_CHECK(a->N % ((a->q_per_kv + 2) * 64) == 0, "dh_gemm: N is not a whole number of query groups");
      return dispatch_bn<DH_EPI_QKV_ROPE>(a, stream);
    case DH_EPI_SWIGLU:
      DH_CHECK(a->out2, "dh_gemm: swiglu epilogue needs out2");
      return dispatch_bn<DH_EPI_SWIGLU>(a, stream);
    case DH_EPI_SWIGLU_BWD:
      DH_CHECK(a->aux && a->out, "dh_gemm: swiglu-bwd epilogue needs aux (h12) and out");
      return dispatch_bn<DH_EPI_SWIGLU_BWD>(a, stream);
    default:
      DH_CHECK(false, "dh_gemm: unknown epilogue %d", a->epilogue);
  }
  return -1;
}

// ------------------------------------------------------------------------------------------------
// dh_gemm_swiglu_streamk_noalloc: the stream-K SwiGLU decode GEMM for callers that may not allocate (no counterpart in the
// reference).  The partial tiles and arrival counters live in a file-scope device
// array that every GPU loading the library holds until exit.  In HOST code the array's name is only a stand-in for the symbol:
// the device address comes from cudaGetSymbolAddress, looked up once per device with that device current.
// ------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_noalloc_scratch[DH_NOALLOC_SCRATCH_BYTES];

namespace dh {
constexpr int NOALLOC_MAX_DEVICES = 64;
static void* g_noalloc_ptr[NOALLOC_MAX_DEVICES];   // device address of g_noalloc_scratch on each device (0 = not looked up yet)

static int noalloc_scratch(void** out) {
  int dev = -1;
  DH_CUDA(cudaGetDevice(&dev));
  DH_CHECK(dev >= 0 && dev < NOALLOC_MAX_DEVICES, "dh_gemm_swiglu_streamk_noalloc: device ordinal %d out of range", dev);
  void* p = __atomic_load_n(&g_noalloc_ptr[dev], __ATOMIC_ACQUIRE);
  if (!p) {
    DH_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));   // same answer from every thread: a benign race
    DH_CHECK(p != nullptr && ((uintptr_t)p & 255) == 0, "dh_gemm_swiglu_streamk_noalloc: bad scratch address %p", p);
    __atomic_store_n(&g_noalloc_ptr[dev], p, __ATOMIC_RELEASE);
  }
  *out = p;
  return 0;
}
}  // namespace dh

extern "C" int dh_gemm_swiglu_streamk_noalloc(const dh_gemm_args* args, void* stream_) {
  using namespace dh;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DH_CHECK(args != nullptr, "dh_gemm_swiglu_streamk_noalloc: null args");
  dh_gemm_args local = *args;
  const int b_layout = local.b_layout & ~DH_B_FROZEN;
  DH_CHECK(local.epilogue == DH_EPI_SWIGLU, "dh_gemm_swiglu_streamk_noalloc: epilogue %d, only DH_EPI_SWIGLU has a stream-K kernel", local.epilogue);
  DH_CHECK(local.M > 0 && local.M <= 2 * BM && local.N > 0 && local.N % SK_BN == 0 && local.K1 > 0 && local.K1 % BK == 0,
           "dh_gemm_swiglu_streamk_noalloc: shape M=%d N=%d K1=%d (needs M <= %d, N %% %d == 0, K1 %% %d == 0)", local.M, local.N,
           local.K1, 2 * BM, SK_BN, BK);
  DH_CHECK(local.K2 == 0 && local.a_layout == 0 && b_layout == 0 && !local.m_limit,
           "dh_gemm_swiglu_streamk_noalloc: K-major operands, no second source, no m_limit");
  DH_CHECK(local.N / SK_BN < sm_count(), "dh_gemm_swiglu_streamk_noalloc: N=%d already fills the %d SMs with whole tiles; use dh_gemm",
           local.N, sm_count());
  const size_t need = streamk_layout(local.M, local.N, local.K1, nullptr);
  DH_CHECK(need <= DH_NOALLOC_SCRATCH_BYTES,
           "dh_gemm_swiglu_streamk_noalloc: M=%d N=%d K1=%d needs %zu bytes of scratch, the static array holds %zu "
           "(DH_NOALLOC_SCRATCH_BYTES); use dh_gemm with a workspace of your own",
           local.M, local.N, local.K1, need, (size_t)DH_NOALLOC_SCRATCH_BYTES);
  void* scratch = nullptr;
  if (noalloc_scratch(&scratch)) return -1;
  // the array keeps the last call's contents: the partial tiles are fully rewritten before they are read, the arrival counters
  // must start at zero (an aborted launch, or a call of another shape, may have left them anywhere)
  const size_t counter_bytes = ((size_t)(local.N / SK_BN) * sizeof(unsigned int) + 255) & ~(size_t)255;
  DH_CUDA(cudaMemsetAsync(scratch, 0, counter_bytes, stream));
  local.workspace = scratch;
  local.workspace_bytes = DH_NOALLOC_SCRATCH_BYTES;
  return dh_gemm(&local, stream_);
}
