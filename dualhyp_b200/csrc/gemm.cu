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
};

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
      a[j] = (h1 / (1.f + __expf(-h1))) * h2;
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
        float sig = 1.f / (1.f + __expf(-h1[j]));
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

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.tiles_m * p.tiles_n;
  const int nkb1 = (p.K1 + BK - 1) / BK;   // a ragged K tail is zero-filled by TMA
  const int nkb = nkb1 + (p.K2 + BK - 1) / BK;

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
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 && lane == 0) {
    // ===================== TMA producer =====================
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m0 = (tile % p.tiles_m) * BM;
      const int n0 = (tile / p.tiles_m) * BN;
      for (int kb = 0; kb < nkb; ++kb, ++it) {
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
  } else if (warp == 1 && lane == 0) {
    // ===================== MMA issuer =====================
    uint32_t it = 0, tcount = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tcount) {
      const int acc = tcount & 1;
      mbar_wait(&tempty_bar[acc], ((tcount >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * BN;
      for (int kb = 0; kb < nkb; ++kb, ++it) {
        const int s = it % STAGES;
        mbar_wait(&full_bar[s], (it / STAGES) & 1);
        tc_fence_after();
        const bool second = kb >= nkb1;
        const int a_mn = second ? p.a2_mn : p.a_mn;
        const int b_mn = second ? p.b2_mn : p.b_mn;
        const uint32_t idesc = umma_idesc_bf16(BM, BN, a_mn, b_mn);
        const uint32_t sa = smem_u32(smem + s * Cfg::STAGE_BYTES);
        const uint32_t sb = sa + Cfg::A_BYTES;
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {
          // K-major: 16 k-elements = 32 bytes inside the 128B swizzle row.  MN-major: 16 k-rows = two
          // 8-row atoms of 1024 B; 64-wide MN groups are 8 KB apart (LBO), 8-row atoms 1 KB apart (SBO).
          const uint64_t da = a_mn ? umma_desc_sw128(sa + k * 2048, 8192, 1024) : umma_desc_sw128(sa + k * 32, 16, 1024);
          const uint64_t db = b_mn ? umma_desc_sw128(sb + k * 2048, 8192, 1024) : umma_desc_sw128(sb + k * 32, 16, 1024);
          tc_mma_ss(d_tmem, da, db, idesc, (kb | k) != 0);
        }
        tc_commit(&empty_bar[s]);   // frees the smem slot when these MMAs have drained
      }
      tc_commit(&tfull_bar[acc]);   // accumulator complete
    }
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    const int q = warp & 3;
    uint32_t tcount = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tcount) {
      const int acc = tcount & 1;
      const int m0 = (tile % p.tiles_m) * BM;
      const int n0 = (tile / p.tiles_m) * BN;
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
        if (row < p.M) epilogue_chunk<EPI>(p, row, n0 + c * 64, v);
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
// Transposing epilogue of the CTA-pair kernel.  A warp owns 32 accumulator rows (TMEM lanes) and handles 64
// accumulator columns per unit: row-wise math in registers (one thread = one row: RoPE / SwiGLU pairs live in
// the same thread), the result staged in a warp-private swizzled shared-memory slab, read back transposed so
// that every global load/store instruction of the warp touches whole 128-byte lines (4 rows x 128 B or
// 2 rows x 256 B) instead of 32 different rows.  Element-wise work against another [M, N] tensor (residual
// add, accumulate, SwiGLU backward against h12) happens on the transposed side with coalesced reads.
// ------------------------------------------------------------------------------------------------
constexpr int STG_BYTES = 8192;   // per epilogue warp: 32 rows x 256 B

// 16-byte chunk c of row r in a slab with 128-byte rows (8 chunks) / 256-byte rows (16 chunks): XOR swizzle inside each
// 128-byte segment keeps both the row-per-lane writes and the transposed reads bank-conflict free.
__device__ __forceinline__ uint32_t stg_off128(int r, int c) { return (uint32_t)(r * 128 + (((c ^ r) & 7) << 4)); }
__device__ __forceinline__ uint32_t stg_off256(int r, int c) { return (uint32_t)(r * 256 + ((c & 8) << 4) + (((c ^ r) & 7) << 4)); }
// 64-byte rows (4 chunks): two rows share a 128-byte bank row
__device__ __forceinline__ uint32_t stg_off64(int r, int c) { return (uint32_t)(r * 64 + (((c ^ (r >> 1)) & 3) << 4)); }

__device__ __forceinline__ void stage_bf16x64(uint8_t* stg, int lane, const float (&f)[64]) {
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    uint4 u;
    u.x = pack_bf16x2(f[8 * c + 0], f[8 * c + 1]); u.y = pack_bf16x2(f[8 * c + 2], f[8 * c + 3]);
    u.z = pack_bf16x2(f[8 * c + 4], f[8 * c + 5]); u.w = pack_bf16x2(f[8 * c + 6], f[8 * c + 7]);
    *reinterpret_cast<uint4*>(stg + stg_off128(lane, c)) = u;
  }
}
// transposed copy-out of a staged [32 x 64] bf16 slab: 4 rows x 128 B per instruction
__device__ __forceinline__ void copy_out_bf16x64(const uint8_t* stg, int lane, bf16* out, long long ld, int row0, int col0, int M) {
  const int rr = lane >> 3, c = lane & 7;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int r = 4 * i + rr;
    const uint4 u = *reinterpret_cast<const uint4*>(stg + stg_off128(r, c));
    if (row0 + r < M) *reinterpret_cast<uint4*>(out + (long long)(row0 + r) * ld + col0 + 8 * c) = u;
  }
}

struct RopeRegs { uint4 c[4], s[4]; };   // cos/sin[pos][0..32) as packed bf16 (table halves are identical)

template <int EPI>
__device__ __forceinline__ void epi_unit(const GemmParams& p, uint8_t* stg, int lane, int row0, int col0, float (&v)[64],
                                         const RopeRegs& rope) {
  const int M = p.M;
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
    stage_bf16x64(stg, lane, v);
    __syncwarp();
    copy_out_bf16x64(stg, lane, reinterpret_cast<bf16*>(p.out), p.ldo, row0, col0, M);
  } else if constexpr (EPI == DH_EPI_F32) {
#pragma unroll
    for (int c = 0; c < 16; ++c)
      *reinterpret_cast<float4*>(stg + stg_off256(lane, c)) =
          make_float4(v[4 * c] * p.alpha, v[4 * c + 1] * p.alpha, v[4 * c + 2] * p.alpha, v[4 * c + 3] * p.alpha);
    __syncwarp();
    const int rr = lane >> 4, c = lane & 15;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int r = 2 * i + rr;
      float4 a = *reinterpret_cast<const float4*>(stg + stg_off256(r, c));
      if (row0 + r < M) {
        float4* o = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + (long long)(row0 + r) * p.ldo + col0 + 4 * c);
        if (p.accumulate) {
          const float4 old = *o;
          a.x += old.x; a.y += old.y; a.z += old.z; a.w += old.w;
        }
        *o = a;
      }
    }
  } else if constexpr (EPI == DH_EPI_RESID_F32) {
    stage_bf16x64(stg, lane, v);   // bf16(acc) is what the reference adds to the fp32 stream
    __syncwarp();
    const int rr = lane >> 4, c = lane & 15;   // output chunk c = 4 floats = 8 staged bytes
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int r = 2 * i + rr;
      const uint2 u = *reinterpret_cast<const uint2*>(stg + stg_off128(r, c >> 1) + ((c & 1) << 3));
      if (row0 + r < M) {
        const long long off = (long long)(row0 + r);
        float4 a = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.aux) + off * p.ldaux + col0 + 4 * c);
        const float2 lo = unpack_bf16x2(u.x), hi = unpack_bf16x2(u.y);
        a.x += lo.x; a.y += lo.y; a.z += hi.x; a.w += hi.y;
        *reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + off * p.ldo + col0 + 4 * c) = a;
      }
    }
  } else if constexpr (EPI == DH_EPI_RESID_BF16) {
    stage_bf16x64(stg, lane, v);
    __syncwarp();
    const int rr = lane >> 3, c = lane & 7;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r = 4 * i + rr;
      const uint4 u = *reinterpret_cast<const uint4*>(stg + stg_off128(r, c));
      if (row0 + r < M) {
        const long long off = (long long)(row0 + r);
        const uint4 x = *reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(p.aux) + off * p.ldaux + col0 + 8 * c);
        const float2 a0 = unpack_bf16x2(u.x), a1 = unpack_bf16x2(u.y), a2 = unpack_bf16x2(u.z), a3 = unpack_bf16x2(u.w);
        const float2 x0 = unpack_bf16x2(x.x), x1 = unpack_bf16x2(x.y), x2 = unpack_bf16x2(x.z), x3 = unpack_bf16x2(x.w);
        uint4 o;
        o.x = pack_bf16x2(x0.x + a0.x, x0.y + a0.y); o.y = pack_bf16x2(x1.x + a1.x, x1.y + a1.y);
        o.z = pack_bf16x2(x2.x + a2.x, x2.y + a2.y); o.w = pack_bf16x2(x3.x + a3.x, x3.y + a3.y);
        *reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(p.out) + off * p.ldo + col0 + 8 * c) = o;
      }
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
          const float a0 = bf16_round(v[j]), b0 = bf16_round(v[j + 32]);
          const float a1 = bf16_round(v[j + 1]), b1 = bf16_round(v[j + 33]);
          v[j] = a0 * cc.x - b0 * ss.x;      v[j + 32] = b0 * cc.x + a0 * ss.x;
          v[j + 1] = a1 * cc.y - b1 * ss.y;  v[j + 33] = b1 * cc.y + a1 * ss.y;
        }
      }
    }
    stage_bf16x64(stg, lane, v);
    __syncwarp();
    copy_out_bf16x64(stg, lane, reinterpret_cast<bf16*>(p.out), p.ldo, row0, col0, M);
  } else if constexpr (EPI == DH_EPI_SWIGLU) {
    // raw h12 (64 cols) in the first 4 KB of the slab, act = silu(h1) * h2 (32 cols, 64-byte rows) in the next 2 KB
    uint8_t* stg2 = stg + 4096;
    if (p.out) stage_bf16x64(stg, lane, v);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float a[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float h1 = bf16_round(v[8 * c + k]), h2 = bf16_round(v[8 * c + k + 32]);
        a[k] = (h1 / (1.f + __expf(-h1))) * h2;
      }
      uint4 u;
      u.x = pack_bf16x2(a[0], a[1]); u.y = pack_bf16x2(a[2], a[3]); u.z = pack_bf16x2(a[4], a[5]); u.w = pack_bf16x2(a[6], a[7]);
      *reinterpret_cast<uint4*>(stg2 + stg_off64(lane, c)) = u;
    }
    __syncwarp();
    if (p.out) copy_out_bf16x64(stg, lane, reinterpret_cast<bf16*>(p.out), p.ldo, row0, col0, M);
    const int rr = lane >> 2, c = lane & 3;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = 8 * i + rr;
      const uint4 u = *reinterpret_cast<const uint4*>(stg2 + stg_off64(r, c));
      if (row0 + r < M) *reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(p.out2) + (long long)(row0 + r) * p.ldo2 + (col0 >> 1) + 8 * c) = u;
    }
  } else if constexpr (EPI == DH_EPI_SWIGLU_BWD) {
    // acc = d(act) columns col0..col0+63; h12 / d(h12) columns 2*col0 .. 2*col0+127 as two [h1 x32 | h2 x32] blocks
    stage_bf16x64(stg, lane, v);
    __syncwarp();
    const int rr = lane >> 3, t = lane & 7, g = t >> 2, cc = t & 3;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r = 4 * i + rr;
      const uint4 du = *reinterpret_cast<const uint4*>(stg + stg_off128(r, g * 4 + cc));
      if (row0 + r < M) {
        const long long off = (long long)(row0 + r);
        const bf16* h = reinterpret_cast<const bf16*>(p.aux) + off * p.ldaux + 2 * col0 + 64 * g + 8 * cc;
        bf16* o = reinterpret_cast<bf16*>(p.out) + off * p.ldo + 2 * col0 + 64 * g + 8 * cc;
        const uint4 h1u = *reinterpret_cast<const uint4*>(h);
        const uint4 h2u = *reinterpret_cast<const uint4*>(h + 32);
        float da[8], h1[8], h2[8], d1[8], d2[8];
        unpack_bf16x8(du, da);
        unpack_bf16x8(h1u, h1);
        unpack_bf16x8(h2u, h2);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float sig = 1.f / (1.f + __expf(-h1[k]));
          d1[k] = da[k] * h2[k] * (sig * (1.f + h1[k] * (1.f - sig)));
          d2[k] = da[k] * (h1[k] * sig);
        }
        *reinterpret_cast<uint4*>(o) = pack_bf16x8(d1);
        *reinterpret_cast<uint4*>(o + 32) = pack_bf16x8(d2);
      }
    }
  }
  __syncwarp();   // the slab is rewritten by the next unit
}

// ------------------------------------------------------------------------------------------------
// CTA-pair kernel (cluster of 2, tcgen05 cta_group::2): one 256 x 256 output tile per pair.  CTA r of the
// pair stages A rows [128 r, 128 r + 128) and B rows (= output columns) [128 r, 128 r + 128) of the tile
// and owns accumulator rows [128 r, +128) x 256 columns in its own TMEM; the leader's MMA thread drives
// both tensor cores, so every B byte is fetched into shared memory once per pair instead of once per SM.
// K-major operands only (dgrads use pre-transposed frozen weights).
// ------------------------------------------------------------------------------------------------
struct PairCfg {
  static constexpr int BN = 256;
  static constexpr int A_BYTES = 128 * BK * 2;
  static constexpr int B_BYTES = 128 * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = 6;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 4 * STG_BYTES + 1024 + 256;
};

template <int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2, const GemmParams p) {
  using Cfg = PairCfg;
  constexpr int STAGES = Cfg::STAGES;
  constexpr int BN = Cfg::BN;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stg_base = smem + STAGES * Cfg::STAGE_BYTES;   // 4 epilogue warps x 8 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(stg_base + 4 * STG_BYTES);
  uint64_t* full_bar = bars;                     // [STAGES]  TMA (both CTAs) -> MMA       (leader's copy is used)
  uint64_t* empty_bar = bars + STAGES;           // [STAGES]  MMA -> TMA                   (multicast to both)
  uint64_t* tfull_bar = bars + 2 * STAGES;       // [2]       MMA -> epilogue              (multicast to both)
  uint64_t* tempty_bar = bars + 2 * STAGES + 2;  // [2]       epilogue (both CTAs) -> MMA  (leader's copy is used)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int pair = blockIdx.x >> 1;
  const int npairs = gridDim.x >> 1;
  const int num_tiles = p.tiles_m * p.tiles_n;
  const int nkb1 = (p.K1 + BK - 1) / BK;
  const int nkb = nkb1 + (p.K2 + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (p.K2 > 0) { tma_prefetch_desc(&tmA2); tma_prefetch_desc(&tmB2); }
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 8); }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc2(tmem_slot, 512);
    tmem_relinquish2();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 && lane == 0) {
    // ===================== TMA producer (each CTA loads its own halves) =====================
    uint32_t it = 0;
    for (int tile = pair; tile < num_tiles; tile += npairs) {
      const int m0 = (tile % p.tiles_m) * 256 + (int)rank * 128;
      const int n0 = (tile / p.tiles_m) * BN + (int)rank * 128;
      for (int kb = 0; kb < nkb; ++kb, ++it) {
        const int s = it % STAGES;
        mbar_wait(&empty_bar[s], ((it / STAGES) & 1) ^ 1);
        uint8_t* sa = smem + s * Cfg::STAGE_BYTES;
        uint8_t* sb = sa + Cfg::A_BYTES;
        if (rank == 0) mbar_expect_tx(&full_bar[s], 2 * Cfg::STAGE_BYTES);
        const uint32_t bar = mapa_shared(smem_u32(&full_bar[s]), 0);
        const bool second = kb >= nkb1;
        const int k0 = (second ? kb - nkb1 : kb) * BK;
        tma_load_2d_pair(sa, second ? &tmA2 : &tmA, k0, m0, bar);
        tma_load_2d_pair(sb, second ? &tmB2 : &tmB, k0, n0, bar);
      }
    }
  } else if (warp == 1 && lane == 0 && rank == 0) {
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
#pragma unroll
        for (int k = 0; k < BK / 16; ++k)
          tc_mma_ss_pair(d_tmem, umma_desc_sw128(sa + k * 32, 16, 1024), umma_desc_sw128(sb + k * 32, 16, 1024), idesc, (kb | k) != 0);
        tc_commit_pair(&empty_bar[s], 3);
      }
      tc_commit_pair(&tfull_bar[acc], 3);
    }
  } else if (warp >= 4) {
    // ===================== epilogue (each CTA drains its own 128 accumulator rows) =====================
    const int q = warp & 3;
    uint32_t tcount = 0;
    for (int tile = pair; tile < num_tiles; tile += npairs, ++tcount) {
      const int acc = tcount & 1;
      const int m0 = (tile % p.tiles_m) * 256 + (int)rank * 128;
      const int n0 = (tile / p.tiles_m) * BN;
      mbar_wait(&tfull_bar[acc], (tcount >> 1) & 1);
      tc_fence_after();
      const int row0 = m0 + q * 32;
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN;
      RopeRegs rope;
      if constexpr (EPI == DH_EPI_QKV_ROPE) {
        const int row = row0 + lane < p.M ? row0 + lane : p.M - 1;
        const int pos = p.positions ? __ldg(p.positions + row) : (row % p.seq_len);
        const uint4* c4 = reinterpret_cast<const uint4*>(p.rope_cos + (long long)pos * 64);
        const uint4* s4 = reinterpret_cast<const uint4*>(p.rope_sin + (long long)pos * 64);
#pragma unroll
        for (int i = 0; i < 4; ++i) { rope.c[i] = __ldg(c4 + i); rope.s[i] = __ldg(s4 + i); }
      }
#pragma unroll 1
      for (int c = 0; c < BN / 64; ++c) {
        uint32_t r0[32], r1[32];
        tmem_ld32(taddr + c * 64, r0);
        tmem_ld32(taddr + c * 64 + 32, r1);
        tmem_wait_ld();
        if (c == BN / 64 - 1) {   // accumulator fully read: hand the TMEM buffer back before the last unit's stores
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(mapa_shared(smem_u32(&tempty_bar[acc]), 0));
        }
        float v[64];
#pragma unroll
        for (int j = 0; j < 32; ++j) { v[j] = __uint_as_float(r0[j]); v[j + 32] = __uint_as_float(r1[j]); }
        epi_unit<EPI>(p, stg_base + q * STG_BYTES, lane, row0, n0 + c * 64, v, rope);
      }
    }
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
  const int tiles = p.tiles_m * p.tiles_n;
  const int grid = tiles < sm_count() ? tiles : sm_count();
  gemm_kernel<BN, EPI><<<grid, GEMM_THREADS, Cfg::SMEM_BYTES, stream>>>(tmA, tmB, tmA2, tmB2, p);
  DH_CUDA(cudaGetLastError());
  return 0;
}

template <int EPI>
static int launch_gemm_pair(const dh_gemm_args* a, cudaStream_t stream) {
  using Cfg = PairCfg;
  static bool attr_set = false;
  if (!attr_set) {
    DH_CUDA(cudaFuncSetAttribute(gemm_pair_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    attr_set = true;
  }
  CUtensorMap tmA, tmB, tmA2, tmB2;
  if (make_operand_map(&tmA, a->A, a->lda, a->M, a->K1, 0, 128)) return -3;
  if (make_operand_map(&tmB, a->B, a->ldb, a->N, a->K1, 0, 128)) return -3;
  if (a->K2 > 0) {
    if (make_operand_map(&tmA2, a->A2, a->lda2, a->M, a->K2, 0, 128)) return -3;
    if (make_operand_map(&tmB2, a->B2, a->ldb2, a->N, a->K2, 0, 128)) return -3;
  } else {
    tmA2 = tmA;
    tmB2 = tmB;
  }
  GemmParams p = make_params(a);
  p.tiles_m = (a->M + 255) / 256;
  p.tiles_n = a->N / Cfg::BN;
  const int tiles = p.tiles_m * p.tiles_n;
  const int pairs = sm_count() / 2;
  const int grid = 2 * (tiles < pairs ? tiles : pairs);
  gemm_pair_kernel<EPI><<<grid, GEMM_THREADS, Cfg::SMEM_BYTES, stream>>>(tmA, tmB, tmA2, tmB2, p);
  DH_CUDA(cudaGetLastError());
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
      a->N % 256 == 0 && a->M >= 1024)
    return launch_gemm_pair<EPI>(a, stream);
  // Tile width: 256 when it divides N and leaves enough tiles to fill the machine, else 128, else 64.
  const long long tiles_m = (a->M + BM - 1) / BM;
  if (a->N % 256 == 0 && tiles_m * (a->N / 256) >= sm_count()) return launch_gemm<256, EPI>(a, stream);
  if (a->N % 128 == 0 && tiles_m * (a->N / 128) >= sm_count() / 2) return launch_gemm<128, EPI>(a, stream);
  return launch_gemm<64, EPI>(a, stream);
}

}  // namespace dh

extern "C" int dh_gemm(const dh_gemm_args* a, void* stream_) {
  using namespace dh;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DH_CHECK(a != nullptr, "dh_gemm: null args");
  DH_CHECK(a->M > 0 && a->N > 0 && a->K1 > 0, "dh_gemm: bad shape M=%d N=%d K1=%d", a->M, a->N, a->K1);
  DH_CHECK(a->N % 64 == 0 && a->K1 % 8 == 0 && a->K2 % 8 == 0, "dh_gemm: N %% 64, K1 %% 8, K2 %% 8 must be 0 (N=%d K1=%d K2=%d)",
           a->N, a->K1, a->K2);
  DH_CHECK(a->A && a->B && (a->out || a->epilogue == DH_EPI_SWIGLU), "dh_gemm: null operand");
  DH_CHECK((a->lda % 8) == 0 && (a->ldb % 8) == 0, "dh_gemm: leading dimensions must be multiples of 8 elements");
  DH_CHECK(((uintptr_t)a->A & 15) == 0 && ((uintptr_t)a->B & 15) == 0, "dh_gemm: operands must be 16-byte aligned");
  if (a->K2 > 0) DH_CHECK(a->A2 && a->B2 && (a->lda2 % 8) == 0 && (a->ldb2 % 8) == 0, "dh_gemm: bad second source");
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
