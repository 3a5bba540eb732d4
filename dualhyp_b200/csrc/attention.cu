// Causal GQA flash attention for head_size 64 on sm_100a with the LLaMA-Adapter gated prefix
// (ger/model.py:270-290, ger/adapter.py:116-146).  tcgen05.mma tiles with TMEM accumulators, operands
// staged by TMA straight out of the group-interleaved QKV row the QKV GEMM wrote (no head split / K,V
// expansion copies: quirk Q4 of SURVEY.md is a broadcast here).  One softmax thread owns one score row
// (TMEM lane), so row max / sum need no shuffles.
//
//   fwd      : CTA = (128 queries, head, batch).  S = Q K^T -> online softmax -> P (bf16, swizzled smem) ->
//              O += P V (V as MN-major B operand).  Prefix term on CUDA cores in the epilogue.
//   bwd prep : delta = rowsum(dO * y) and the whole prefix/gate backward (CUDA cores).
//   bwd dKdV : CTA = (128 keys, kv head, batch), loops over the q heads of the group and q tiles >= diagonal.
//   bwd dQ   : CTA = (128 queries, head, batch), loops over k tiles <= diagonal.
//   RoPE backward (conjugate rotation) is fused into the dQ / dK epilogues.
//   decode   : single-query attention over the KV cache (CUDA cores; HBM-bound).
#include <math.h>

#include <stdlib.h>

#include "common.cuh"
#include "../../include/dualhyp.h"

namespace dh {

#ifdef DH_ATTN_TRACE
#define TRC_DECL long long trc_t[48]; int trc_n = 0; const bool trc_on = (blockIdx.x == 0 && blockIdx.y == 1 && blockIdx.z == 3);
#define TRC(tag) do { if (trc_on && trc_n < 48) { trc_t[trc_n++] = clock64(); } } while (0)
#define TRC_DUMP(who) do { if (trc_on && lane == 0) { for (int i_ = 1; i_ < trc_n; ++i_) printf("%s w%d ev%d +%lld (abs %lld)\n", who, warp, i_, trc_t[i_] - trc_t[i_ - 1], trc_t[i_] - trc_t[0]); } } while (0)
#else
#define TRC_DECL
#define TRC(tag)
#define TRC_DUMP(who)
#endif

constexpr int ATT_BQ = 128;
constexpr int ATT_BK = 128;
constexpr int HD = 64;
constexpr int TILE_BYTES = 128 * 128;          // one [128 x 64] bf16 tile (128 B rows)
constexpr int MAX_PREFIX = 16;
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;

// 32 consecutive K-columns [32c, 32c+32) of row r of a [128 x 128] bf16 A operand kept as two K-major
// [128][64] tiles with the 128B swizzle.
__device__ __forceinline__ void store_a_chunk(uint8_t* base, int r, int c, const float (&f)[32]) {
  uint8_t* rowp = base + (c >> 1) * TILE_BYTES + r * 128;
  const int chunk0 = (c & 1) * 4;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    uint4 u;
    u.x = pack_bf16x2(f[8 * i + 0], f[8 * i + 1]); u.y = pack_bf16x2(f[8 * i + 2], f[8 * i + 3]);
    u.z = pack_bf16x2(f[8 * i + 4], f[8 * i + 5]); u.w = pack_bf16x2(f[8 * i + 6], f[8 * i + 7]);
    *reinterpret_cast<uint4*>(rowp + (((chunk0 + i) ^ (r & 7)) << 4)) = u;
  }
}

// S/dP style MMA: D[128 x 128] = A[128 x 64] * B[128 x 64]^T, both K-major tiles.
__device__ __forceinline__ void mma_qk(uint32_t d_tmem, uint32_t a_smem, uint32_t b_smem) {
  constexpr uint32_t idesc = umma_idesc_bf16(128, 128, 0, 0);
#pragma unroll
  for (int k = 0; k < 4; ++k)
    tc_mma_ss(d_tmem, umma_desc_sw128(a_smem + k * 32, 16, 1024), umma_desc_sw128(b_smem + k * 32, 16, 1024), idesc, k != 0);
}
// PV style MMA: D[128 x 64] (+)= A[128 x 128] * B[128 x 64], A = two K-major tiles, B = [k rows][64 n] (MN-major).
__device__ __forceinline__ void mma_pv(uint32_t d_tmem, uint32_t a_smem, uint32_t b_smem, bool accumulate) {
  constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 0, 1);
#pragma unroll
  for (int k = 0; k < 8; ++k)
    tc_mma_ss(d_tmem, umma_desc_sw128(a_smem + (k >> 2) * TILE_BYTES + (k & 3) * 32, 16, 1024),
              umma_desc_sw128(b_smem + k * 2048, 8192, 1024), idesc, accumulate || k != 0);
}

struct AttnParams {
  int B, T, G, qpk, H, slots;       // slots = G*(qpk+2)
  int prefix_len;
  float scale, scale_log2;
  long long ld_qkv;                 // elements
  const bf16* qkv;
  bf16* out;                        // fwd
  float* lse;
  const bf16* prefix_kv;
  const float* gate;
  // bwd
  const bf16* o;
  const bf16* dout;
  bf16* dqkv;
  float* delta;
  float* d_prefix_kv;
  float* d_gate;
  int need_dqkv;
  const bf16* rope_cos;
  const bf16* rope_sin;
  // forward over packed variable-length sequences (no pad rows): sequence b = rows [cu[b], cu[b + 1]); NULL = B x T
  const int* cu_seqlens;
  long long total_rows;
};

// index of (batch/sequence b, head h, position t inside the sequence) in the lse / delta statistics: [B, H, T] for padded batches,
// [H, total_rows] (row = row_base + t) for packed variable-length sequences
__device__ __forceinline__ long long stat_index(const AttnParams& p, int b, int h, int row_base, int t) {
  return p.cu_seqlens ? (long long)h * p.total_rows + row_base + t : ((long long)b * p.H + h) * p.T + t;
}
// the same for a flat token row m of the whole batch
__device__ __forceinline__ long long stat_index_flat(const AttnParams& p, int h, long long m) {
  return p.cu_seqlens ? (long long)h * p.total_rows + m : ((long long)(m / p.T) * p.H + h) * p.T + (m % p.T);
}

__device__ __forceinline__ void load_prefix_to_smem(const AttnParams& p, int g, float* sak, float* sav) {
  const int n = p.prefix_len * HD;
  const long long koff = (long long)(g * (p.qpk + 2) + p.qpk) * HD;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int a = i / HD, d = i % HD;
    sak[i] = __bfloat162float(p.prefix_kv[(long long)a * p.ld_qkv + koff + d]);
    sav[i] = __bfloat162float(p.prefix_kv[(long long)a * p.ld_qkv + koff + HD + d]);
  }
}

// conjugate (backward) NeoX rotation of a 64-wide fp32 row held as two 32-halves
__device__ __forceinline__ void rope_conj_halves(float (&lo)[32], float (&hi)[32], const bf16* cs, const bf16* sn, int pos) {
  const uint4* c4 = reinterpret_cast<const uint4*>(cs + (long long)pos * 64);
  const uint4* s4 = reinterpret_cast<const uint4*>(sn + (long long)pos * 64);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float c[8], s[8];
    unpack8(c4[i], c);
    unpack8(s4[i], s);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float a = lo[8 * i + j], b = hi[8 * i + j];
      lo[8 * i + j] = a * c[j] + b * s[j];
      hi[8 * i + j] = b * c[j] - a * s[j];
    }
  }
}

// =================================================================================================
// forward.  CTA = (q tile pair {nq-1-x, x}, head, batch): every CTA does nq+1 key tiles, so the grid is
// balanced whatever the causal triangle looks like; two CTAs share an SM (256 TMEM columns each) and
// fill each other's bubbles.  Per key tile:
//   MMA warp   : S = Q K^T (tcgen05, fp32 in TMEM)            -> s_full
//   softmax    : 8 warps; two threads per query row, 64 score columns each, straight into registers
//                (-> s_free: the MMA warp issues the NEXT tile's S while this one is exponentiated);
//                the row max is exchanged through shared memory; lazy rescaling of O (only when the
//                max grew by more than 2^8); P = 2^(s - m) packed to bf16 and written back to TMEM -> p_full
//   MMA warp   : O += P V with P as the TMEM A operand, V the MN-major smem B operand  -> o_done
// K and V tiles are double buffered.  The adapter prefix (ger/adapter.py:129-146) rides the same
// pipes: S_a = Q AK^T is a 16-column MMA issued before the first key tile, its softmax weights stay
// in registers and, once the causal row sum l is known, gate * l * softmax(S_a) is the A operand of
// one last K=16 MMA into O, so (O + l gate P_a AV) / l = causal + gate * prefix.
// =================================================================================================
constexpr int FWD_SM_WARPS = 8;                       // softmax warps
constexpr int FWD_THREADS = (FWD_SM_WARPS + 2) * 32;  // + TMA warp + MMA warp
constexpr int FWD_OFF_K = 2 * TILE_BYTES, FWD_OFF_V = 4 * TILE_BYTES, FWD_OFF_AK = 6 * TILE_BYTES, FWD_OFF_AV = FWD_OFF_AK + 2048,
              FWD_OFF_X = FWD_OFF_AV + 2048 /* float [2 parity][2 halves][128] */, FWD_OFF_PW = FWD_OFF_X + 2048 /* float [16][128] */,
              FWD_OFF_BAR = FWD_OFF_PW + MAX_PREFIX * 128 * 4;
constexpr int FWD_SMEM = FWD_OFF_BAR + 256 + 1024;
constexpr float RESCALE_THRESHOLD = 8.f;   // log2 units

// D[128 x 64] (+)= A[128 x 128] (TMEM, bf16 pairs per column) * B[128 x 64] (smem tile [k rows][64], MN-major)
__device__ __forceinline__ void mma_ts_k128(uint32_t d_tmem, uint32_t a_tmem, uint32_t b_smem, bool accumulate) {
  constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 0, 1);
#pragma unroll
  for (int k = 0; k < 8; ++k)
    tc_mma_ts(d_tmem, a_tmem + k * 8, umma_desc_sw128(b_smem + k * 2048, 8192, 1024), idesc, accumulate || k != 0);
}

// max over one 32-column chunk of a score row; TRI: only columns <= lane are visible (diagonal chunk)
template <bool TRI>
__device__ __forceinline__ void chunk_max(const uint32_t (&s)[32], int lane, float& m0, float& m1) {
#pragma unroll
  for (int i = 0; i < 32; i += 2) {
    float a = __uint_as_float(s[i]), c = __uint_as_float(s[i + 1]);
    if (TRI) { a = (i <= lane) ? a : -INFINITY; c = (i + 1 <= lane) ? c : -INFINITY; }
    m0 = fmaxf(m0, a);
    m1 = fmaxf(m1, c);
  }
}
// P = 2^(s * scale_log2 - m) for one chunk -> 16 packed bf16 pairs; the row sum is taken before rounding
template <bool TRI>
__device__ __forceinline__ void chunk_exp(const uint32_t (&s)[32], int lane, float scale_log2, float neg_m, float& sum0, float& sum1,
                                          uint32_t (&out)[16]) {
#pragma unroll
  for (int i = 0; i < 32; i += 2) {
    float a = fast_exp2(fmaf(__uint_as_float(s[i]), scale_log2, neg_m));
    float c = fast_exp2(fmaf(__uint_as_float(s[i + 1]), scale_log2, neg_m));
    if (TRI) { a = (i <= lane) ? a : 0.f; c = (i + 1 <= lane) ? c : 0.f; }
    sum0 += a;
    sum1 += c;
    out[i >> 1] = pack_bf16x2(a, c);
  }
}
// mode: 0 = chunk fully visible, 1 = triangular (diagonal chunk), 2 = fully masked
__device__ __forceinline__ void chunk_max_mode(int mode, const uint32_t (&s)[32], int lane, float& m0, float& m1) {
  if (mode == 0) chunk_max<false>(s, lane, m0, m1);
  else if (mode == 1) chunk_max<true>(s, lane, m0, m1);
}
__device__ __forceinline__ void chunk_exp_mode(int mode, const uint32_t (&s)[32], int lane, float scale_log2, float neg_m, float& sum0,
                                               float& sum1, uint32_t (&out)[16]) {
  if (mode == 0) chunk_exp<false>(s, lane, scale_log2, neg_m, sum0, sum1, out);
  else if (mode == 1) chunk_exp<true>(s, lane, scale_log2, neg_m, sum0, sum1, out);
  else {
#pragma unroll
    for (int i = 0; i < 16; ++i) out[i] = 0u;
  }
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

__global__ void __launch_bounds__(FWD_THREADS, 2) attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQKV,
                                                                  const __grid_constant__ CUtensorMap tmPFX, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;                    // [2 items]
  uint8_t* sK = smem + FWD_OFF_K;        // [2 stages]
  uint8_t* sV = smem + FWD_OFF_V;        // [2 stages]
  uint8_t* sAK = smem + FWD_OFF_AK;      // [16 x 64] prefix keys (rows >= prefix_len are zero)
  uint8_t* sAV = smem + FWD_OFF_AV;
  float* sX = reinterpret_cast<float*>(smem + FWD_OFF_X);   // row max / row sum exchange between the two halves of a row
  float* sPW = reinterpret_cast<float*>(smem + FWD_OFF_PW);  // gate * softmax over the prefix keys, [a][row]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + FWD_OFF_BAR);
  uint64_t *q_full = bars /*[2]*/, *k_full = bars + 2 /*[2]*/, *k_free = bars + 4 /*[2]*/, *v_full = bars + 6 /*[2]*/,
           *v_free = bars + 8 /*[2]*/, *s_full = bars + 10, *s_free = bars + 11, *p_full = bars + 12, *o_done = bars + 13,
           *pfx_full = bars + 14;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // shfl: warp-uniform for the compiler
  TRC_DECL
  TRC(0);
  const int h = blockIdx.y, b = blockIdx.z;
  // one length per batch (rows b T .. b T + T), or packed sequences of their own lengths starting at cu_seqlens[b]
  const int row_base = p.cu_seqlens ? __ldg(p.cu_seqlens + b) : b * p.T;
  const int T_b = p.cu_seqlens ? __ldg(p.cu_seqlens + b + 1) - row_base : p.T;
  const int nq = (T_b + ATT_BQ - 1) / ATT_BQ;
  if ((int)blockIdx.x >= (nq + 1) / 2) return;      // the grid is sized for the longest sequence
  const int g = h / p.qpk;
  const int q_slot = g * (p.qpk + 2) + (h % p.qpk), k_slot = g * (p.qpk + 2) + p.qpk, v_slot = k_slot + 1;
  const int qi_first = nq - 1 - (int)blockIdx.x, qi_second = (int)blockIdx.x;
  const int n_items = (qi_first == qi_second) ? 1 : 2;
  const bool has_prefix = p.prefix_len > 0;

  if (warp == FWD_SM_WARPS && lane == 0) {
    // the producer thread initialises the barriers and starts the first loads before the CTA-wide sync
    for (int i = 0; i < 2; ++i) {
      mbar_init(&q_full[i], 1); mbar_init(&k_full[i], 1); mbar_init(&k_free[i], 1); mbar_init(&v_full[i], 1); mbar_init(&v_free[i], 1);
    }
    mbar_init(s_full, 1); mbar_init(s_free, FWD_SM_WARPS * 32); mbar_init(p_full, FWD_SM_WARPS * 32); mbar_init(o_done, 1);
    mbar_init(pfx_full, 1);
    fence_barrier_init();
    mbar_expect_tx(&q_full[0], TILE_BYTES);
    tma_load_3d(sQ, &tmQKV, 0, q_slot, row_base + qi_first * ATT_BQ, &q_full[0]);
    if (has_prefix) {
      mbar_expect_tx(pfx_full, 4096);
      tma_load_3d(sAK, &tmPFX, 0, k_slot, 0, pfx_full);
      tma_load_3d(sAV, &tmPFX, 0, v_slot, 0, pfx_full);
    }
    mbar_expect_tx(&k_full[0], TILE_BYTES);
    tma_load_3d(sK, &tmQKV, 0, k_slot, row_base, &k_full[0]);
    mbar_expect_tx(&v_full[0], TILE_BYTES);
    tma_load_3d(sV, &tmQKV, 0, v_slot, row_base, &v_full[0]);
  }
  if (warp == FWD_SM_WARPS + 1) { tmem_alloc(tmem_slot, 256); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  const uint32_t tS = tmem, tO = tmem + 128, tP = tmem + 192;

  if (warp == FWD_SM_WARPS) {
    // ---------------- TMA producer (tile 0 of item 0 and Q of item 0 are already in flight) ----------------
    if (lane == 0) {
      if (n_items == 2) {
        mbar_expect_tx(&q_full[1], TILE_BYTES);
        tma_load_3d(sQ + TILE_BYTES, &tmQKV, 0, q_slot, row_base + qi_second * ATT_BQ, &q_full[1]);
      }
      int n = 0;
      for (int it = 0; it < n_items; ++it) {
        const int qi = it ? qi_second : qi_first;
        for (int j = 0; j <= qi; ++j, ++n) {
          if (n == 0) continue;
          const int st = n & 1;
          if (n >= 2) mbar_wait(&k_free[st], ((n >> 1) - 1) & 1);
          mbar_expect_tx(&k_full[st], TILE_BYTES);
          tma_load_3d(sK + st * TILE_BYTES, &tmQKV, 0, k_slot, row_base + j * ATT_BK, &k_full[st]);
          if (n >= 2) mbar_wait(&v_free[st], ((n >> 1) - 1) & 1);
          mbar_expect_tx(&v_full[st], TILE_BYTES);
          tma_load_3d(sV + st * TILE_BYTES, &tmQKV, 0, v_slot, row_base + j * ATT_BK, &v_full[st]);
        }
      }
    }
  } else if (warp == FWD_SM_WARPS + 1) {
    // ---------------- MMA issuer: the whole warp walks the loop (uniform operands), one elected lane issues ----------------
    int n = 0, es = 0, ep = 0;
    if (has_prefix) mbar_wait(pfx_full, 0);
    for (int it = 0; it < n_items; ++it) {
      const int nkv = (it ? qi_second : qi_first) + 1;
      const uint32_t q_addr = smem_u32(sQ + it * TILE_BYTES);
      mbar_wait(&q_full[it], 0);
      tc_fence_after();
      if (has_prefix) {
        const uint32_t ak_addr = smem_u32(sAK);
        if (elect_one()) {
          constexpr uint32_t idesc16 = umma_idesc_bf16(128, 16, 0, 0);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            tc_mma_ss(tS, umma_desc_sw128(q_addr + k * 32, 16, 1024), umma_desc_sw128(ak_addr + k * 32, 16, 1024), idesc16, k != 0);
          tc_commit(s_full);
        }
        __syncwarp();
        mbar_wait(s_free, es & 1); ++es;
        tc_fence_after();
      }
      mbar_wait(&k_full[n & 1], (n >> 1) & 1);
      tc_fence_after();
      {
        const uint32_t k_addr = smem_u32(sK + (n & 1) * TILE_BYTES);
        if (elect_one()) {
          mma_qk(tS, q_addr, k_addr);
          tc_commit(&k_free[n & 1]);
          tc_commit(s_full);
        }
        __syncwarp();
      }
      for (int j = 0; j < nkv; ++j, ++n) {
        mbar_wait(s_free, es & 1); ++es;           // S(j) is in registers
        tc_fence_after();
        if (j + 1 < nkv) {
          const int n1 = n + 1;
          mbar_wait(&k_full[n1 & 1], (n1 >> 1) & 1);
          tc_fence_after();
          const uint32_t k_addr = smem_u32(sK + (n1 & 1) * TILE_BYTES);
          if (elect_one()) {
            mma_qk(tS, q_addr, k_addr);
            tc_commit(&k_free[n1 & 1]);
            tc_commit(s_full);
          }
          __syncwarp();
        }
        mbar_wait(p_full, ep & 1); ++ep;           // P(j) is in TMEM, O rescaled if needed
        mbar_wait(&v_full[n & 1], (n >> 1) & 1);
        tc_fence_after();
        const uint32_t v_addr = smem_u32(sV + (n & 1) * TILE_BYTES);
        const bool acc = j > 0;
        if (elect_one()) {
          mma_ts_k128(tO, tP, v_addr, acc);
          tc_commit(&v_free[n & 1]);
          tc_commit(o_done);
        }
        __syncwarp();
      }
      if (has_prefix) {
        mbar_wait(p_full, ep & 1); ++ep;
        tc_fence_after();
        const uint32_t av_addr = smem_u32(sAV);
        if (elect_one()) {
          tc_mma_ts(tO, tP, umma_desc_sw128(av_addr, 8192, 1024), umma_idesc_bf16(128, 64, 0, 1), 1);
          tc_commit(o_done);
        }
        __syncwarp();
      }
    }
  } else {
    // ---------------- softmax / epilogue: two threads per query row, 64 score columns each ----------------
    const int quad = warp & 3, hf = warp >> 2;
    const int r = quad * 32 + lane;
    const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
    const int aT = p.prefix_len;
    const float sl2 = p.scale_log2;
    int fs = 0;       // s_full events consumed
    int np = 0;       // p_full arrivals made (one o_done event follows each)
    int wo = 0;       // o_done events waited for
    int xc = 0;       // exchanges through sX (parity selects the buffer)
    for (int it = 0; it < n_items; ++it) {
      const int qi = it ? qi_second : qi_first;
      const int nkv = qi + 1;
      const int tq = qi * ATT_BQ + r;
      if (has_prefix) {
        mbar_wait(s_full, fs & 1); ++fs;
        tc_fence_after();
        uint32_t v[16];
        tmem_ld16(tS + lane_off, v);
        tmem_wait_ld();
        tc_fence_before();
        mbar_arrive(s_free);
        if (hf == 0) {
          float mx = -INFINITY;
#pragma unroll
          for (int a = 0; a < MAX_PREFIX; ++a) if (a < aT) mx = fmaxf(mx, __uint_as_float(v[a]));
          float sum = 0.f;
#pragma unroll
          for (int a = 0; a < MAX_PREFIX; ++a) {
            const float e = (a < aT) ? fast_exp2((__uint_as_float(v[a]) - mx) * sl2) : 0.f;
            v[a] = __float_as_uint(e);
            sum += e;
          }
          const float gs = p.gate[h] / sum;
#pragma unroll
          for (int a = 0; a < MAX_PREFIX; ++a) sPW[a * 128 + r] = __uint_as_float(v[a]) * gs;   // read back by this thread only
        }
      }
      float m_ref = -INFINITY, l = 0.f;
      for (int j = 0; j < nkv; ++j) {
        mbar_wait(s_full, fs & 1); ++fs;
        TRC(1);
        tc_fence_after();
        uint32_t s0[32], s1[32];
        tmem_ld32(tS + lane_off + 64 * hf, s0);
        tmem_ld32(tS + lane_off + 64 * hf + 32, s1);
        tmem_wait_ld();
        TRC(2);
        tc_fence_before();
        mbar_arrive(s_free);
        // diagonal tile: rows of quad w see 32-column chunks c < w unmasked, chunk w triangular, chunks c > w masked
        int mode0 = 0, mode1 = 0;
        if (j == qi) {
          const int c0 = 2 * hf, c1 = 2 * hf + 1;
          mode0 = c0 < quad ? 0 : (c0 == quad ? 1 : 2);
          mode1 = c1 < quad ? 0 : (c1 == quad ? 1 : 2);
        }
        float mx0 = -INFINITY, mx1 = -INFINITY;
        chunk_max_mode(mode0, s0, lane, mx0, mx1);
        chunk_max_mode(mode1, s1, lane, mx0, mx1);
        float* xb = sX + (xc & 1) * 256; ++xc;
        xb[hf * 128 + r] = fmaxf(mx0, mx1);
        named_bar_sync(1, FWD_SM_WARPS * 32);
        const float m_new = fmaxf(m_ref, fmaxf(fmaxf(mx0, mx1), xb[(hf ^ 1) * 128 + r]) * sl2);
        TRC(3);
        // lazy rescaling: keep the stale reference max while the true max is within 2^8 of it
        if (__any_sync(0xffffffffu, m_new - m_ref > RESCALE_THRESHOLD)) {
          const float alpha = fast_exp2(m_ref - m_new);     // 0 on the first tile (m_ref = -inf)
          l *= alpha;
          m_ref = m_new;
          if (j > 0) {
            while (wo < np) { mbar_wait(o_done, wo & 1); ++wo; }
            tc_fence_after();
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              uint32_t o[16];
              tmem_ld16(tO + lane_off + 32 * hf + c * 16, o);
              tmem_wait_ld();
#pragma unroll
              for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
              tmem_st16(tO + lane_off + 32 * hf + c * 16, o);
            }
          }
        }
        while (wo < np) { mbar_wait(o_done, wo & 1); ++wo; }   // P V(j-1) has read the previous P
        TRC(4);
        tc_fence_after();
        const float neg_m = -m_ref;
        float sum0 = 0.f, sum1 = 0.f;
        uint32_t pk[16];
        chunk_exp_mode(mode0, s0, lane, sl2, neg_m, sum0, sum1, pk);
        tmem_st16(tP + lane_off + 32 * hf, pk);
        chunk_exp_mode(mode1, s1, lane, sl2, neg_m, sum0, sum1, pk);
        tmem_st16(tP + lane_off + 32 * hf + 16, pk);
        l += sum0 + sum1;
        TRC(5);
        tmem_wait_st();
        TRC(6);
        tc_fence_before();
        mbar_arrive(p_full); ++np;
      }
      // total row sum: exchange the two halves
      {
        float* xb = sX + (xc & 1) * 256; ++xc;
        xb[hf * 128 + r] = l;
        named_bar_sync(1, FWD_SM_WARPS * 32);
        l += xb[(hf ^ 1) * 128 + r];
      }
      if (has_prefix) {
        // A operand of the last MMA: l * gate * softmax_a, so that the common 1/l below leaves gate * prefix
        while (wo < np) { mbar_wait(o_done, wo & 1); ++wo; }
        tc_fence_after();
        if (hf == 0) {
          uint32_t pk[8];
#pragma unroll
          for (int a = 0; a < 8; ++a) pk[a] = pack_bf16x2(sPW[(2 * a) * 128 + r] * l, sPW[(2 * a + 1) * 128 + r] * l);
          tmem_st8(tP + lane_off, pk);
          tmem_wait_st();
        }
        tc_fence_before();
        mbar_arrive(p_full); ++np;
      }
      while (wo < np) { mbar_wait(o_done, wo & 1); ++wo; }
      TRC(7);
      tc_fence_after();
      const float inv_l = 1.f / l;
      // TMEM loads are .sync.aligned: every lane issues them; only the global stores are predicated.
      const bool valid = tq < T_b;
      bf16* orow = p.out + (long long)(row_base + (valid ? tq : 0)) * (p.H * HD) + h * HD + 32 * hf;
      uint32_t o[32];
      tmem_ld32(tO + lane_off + 32 * hf, o);
      tmem_wait_ld();
      tc_fence_before();
      if (valid) {
        uint4* o4 = reinterpret_cast<uint4*>(orow);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint4 u;
          u.x = pack_bf16x2(__uint_as_float(o[8 * i + 0]) * inv_l, __uint_as_float(o[8 * i + 1]) * inv_l);
          u.y = pack_bf16x2(__uint_as_float(o[8 * i + 2]) * inv_l, __uint_as_float(o[8 * i + 3]) * inv_l);
          u.z = pack_bf16x2(__uint_as_float(o[8 * i + 4]) * inv_l, __uint_as_float(o[8 * i + 5]) * inv_l);
          u.w = pack_bf16x2(__uint_as_float(o[8 * i + 6]) * inv_l, __uint_as_float(o[8 * i + 7]) * inv_l);
          o4[i] = u;
        }
        if (hf == 0) {
          p.lse[stat_index(p, b, h, row_base, tq)] = (m_ref + log2f(l)) * LN2;
        }
      }
      TRC(8);
    }
  }
  if (warp == 0) TRC_DUMP("sm");
  tc_fence_before();
  __syncthreads();
  if (warp == FWD_SM_WARPS + 1) { tc_fence_after(); tmem_dealloc(tmem, 256); }
}

// =================================================================================================
// backward dK/dV.  CTA = (k tile pair {nq-1-x, x}, kv group, batch), one CTA per SM, flat loop over the
// qpk heads of the group and the q tiles i >= j of each key tile j (every CTA does qpk*(nq+1) iterations).
// The key row is the TMEM lane: S^T = K Q^T and dP^T = V dO^T (SS MMAs) land in TMEM, eight softmax warps
// (four threads per key row, 32 query columns each) pull them into registers (-> st_free: the MMA warp issues
// the NEXT iteration's S^T / dP^T right away), form P^T = 2^(s - lse) and dS^T = P^T (dP^T - delta), and
// write both back to TMEM as bf16 A operands of dV += P^T dO and dK += dS^T Q (TS MMAs).  dV / dK stay in
// TMEM over all heads of the group (the GQA reduction is free).  Q / dO tiles (+ their lse / delta rows) are
// double buffered; the K/V tiles of the two key tiles of the pair have a buffer each.
// =================================================================================================
constexpr int DKV_SM_WARPS = 16;
constexpr int DKV_STAGES = 3;     // Q / dO ring
constexpr int DKV_THREADS = (DKV_SM_WARPS + 2) * 32;
constexpr int DKV_OFF_QDO = 4 * TILE_BYTES, DKV_OFF_STAT = (4 + 2 * DKV_STAGES) * TILE_BYTES, DKV_OFF_BAR = DKV_OFF_STAT + DKV_STAGES * 256 * 4;
constexpr int DKV_SMEM = DKV_OFF_BAR + 256 + 1024;

__device__ __forceinline__ float4 lds128(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
// 16 columns [c0, c0+16) of the backward elementwise step: P = 2^(s*sl2 - lse2), dS = P (dP - delta), packed bf16 pairs.
// MODE 0: all visible, 1: columns <= lane visible, 2: columns >= lane visible, 3: nothing visible (c = column index
// inside the 32-column chunk).  lse2 / delta come per column (COLSTAT: shared-memory address of the 16 floats of this
// piece, delta 512 bytes further) or per row.
template <int MODE, bool COLSTAT, bool WANT_P>
__device__ __forceinline__ void bwd_piece(const uint32_t* s, const uint32_t* dp, int c0, int lane, float sl2, uint32_t stat_addr, float lse2r,
                                          float deltar, uint32_t (&pp)[8], uint32_t (&ds)[8]) {
  if (MODE == 3) {
#pragma unroll
    for (int i = 0; i < 8; ++i) { if (WANT_P) pp[i] = 0u; ds[i] = 0u; }
    return;
  }
#pragma unroll
  for (int i4 = 0; i4 < 4; ++i4) {
    float l4[4], d4[4];
    if (COLSTAT) {
      const float4 a = lds128(stat_addr + 16 * i4);
      const float4 c = lds128(stat_addr + 512 + 16 * i4);
      l4[0] = a.x; l4[1] = a.y; l4[2] = a.z; l4[3] = a.w;
      d4[0] = c.x; d4[1] = c.y; d4[2] = c.z; d4[3] = c.w;
    } else {
#pragma unroll
      for (int k = 0; k < 4; ++k) { l4[k] = lse2r; d4[k] = deltar; }
    }
    float pr[4], dv[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int i = 4 * i4 + k;
      float e = fast_exp2(fmaf(__uint_as_float(s[i]), sl2, -l4[k]));
      if (MODE == 1) e = (c0 + i <= lane) ? e : 0.f;
      if (MODE == 2) e = (c0 + i >= lane) ? e : 0.f;
      pr[k] = e;
      dv[k] = e * (__uint_as_float(dp[i]) - d4[k]);
    }
    if (WANT_P) { pp[2 * i4] = pack_bf16x2(pr[0], pr[1]); pp[2 * i4 + 1] = pack_bf16x2(pr[2], pr[3]); }
    ds[2 * i4] = pack_bf16x2(dv[0], dv[1]);
    ds[2 * i4 + 1] = pack_bf16x2(dv[2], dv[3]);
  }
}
template <bool COLSTAT, bool WANT_P>
__device__ __forceinline__ void bwd_piece_mode(int mode, const uint32_t* s, const uint32_t* dp, int c0, int lane, float sl2,
                                               uint32_t stat_addr, float lse2r, float deltar, uint32_t (&pp)[8], uint32_t (&ds)[8]) {
  if (mode == 0) bwd_piece<0, COLSTAT, WANT_P>(s, dp, c0, lane, sl2, stat_addr, lse2r, deltar, pp, ds);
  else if (mode == 1) bwd_piece<1, COLSTAT, WANT_P>(s, dp, c0, lane, sl2, stat_addr, lse2r, deltar, pp, ds);
  else if (mode == 2) bwd_piece<2, COLSTAT, WANT_P>(s, dp, c0, lane, sl2, stat_addr, lse2r, deltar, pp, ds);
  else bwd_piece<3, COLSTAT, WANT_P>(s, dp, c0, lane, sl2, stat_addr, lse2r, deltar, pp, ds);
}

// 8 + 8 accumulator columns (d in [8qt, 8qt+8) and the matching d + 32) -> bf16 row segments in global memory,
// optionally through the conjugate RoPE rotation at position `pos`.
__device__ __forceinline__ void store_rot_pairs(bf16* row, int qt, uint32_t (&lo)[8], uint32_t (&hi)[8], float scale, const bf16* cs,
                                                const bf16* sn, int pos) {
  float a[8], c[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { a[i] = __uint_as_float(lo[i]) * scale; c[i] = __uint_as_float(hi[i]) * scale; }
  if (cs) {
    float cc[8], ss[8];
    unpack8(*reinterpret_cast<const uint4*>(cs + (long long)pos * 64 + 8 * qt), cc);
    unpack8(*reinterpret_cast<const uint4*>(sn + (long long)pos * 64 + 8 * qt), ss);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float x = a[j], y = c[j];
      a[j] = x * cc[j] + y * ss[j];
      c[j] = y * cc[j] - x * ss[j];
    }
  }
  *reinterpret_cast<uint4*>(row + 8 * qt) = pack8(a);
  *reinterpret_cast<uint4*>(row + 32 + 8 * qt) = pack8(c);
}

__global__ void __launch_bounds__(DKV_THREADS, 1) attn_bwd_dkv_kernel(const __grid_constant__ CUtensorMap tmQKV,
                                                                       const __grid_constant__ CUtensorMap tmDO, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sKV = smem;                       // [2 items][K | V]
  uint8_t* sQdO = smem + DKV_OFF_QDO;        // [DKV_STAGES][Q | dO]
  float* sStat = reinterpret_cast<float*>(smem + DKV_OFF_STAT);   // [DKV_STAGES][lse*log2e 128 | delta 128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + DKV_OFF_BAR);
  uint64_t *kv_full = bars /*[2]*/, *qdo_full = bars + 2 /*[STAGES]*/, *qdo_free = bars + 2 + DKV_STAGES /*[STAGES]*/,
           *st_full = bars + 2 + 2 * DKV_STAGES, *st_free = st_full + 1, *pds_full = st_full + 2, *acc_done = st_full + 3;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(st_full + 4);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // shfl: the compiler sees a warp-uniform value
  TRC_DECL
  TRC(0);
  const int g = blockIdx.y, b = blockIdx.z;
  // one length per batch (rows b T .. b T + T), or packed sequences of their own lengths starting at cu_seqlens[b]
  const int row_base = p.cu_seqlens ? __ldg(p.cu_seqlens + b) : b * p.T;
  const int T_b = p.cu_seqlens ? __ldg(p.cu_seqlens + b + 1) - row_base : p.T;
  const int nq = (T_b + ATT_BQ - 1) / ATT_BQ;
  if ((int)blockIdx.x >= (nq + 1) / 2) return;      // the grid is sized for the longest sequence
  const int k_slot = g * (p.qpk + 2) + p.qpk, v_slot = k_slot + 1;
  const int j_first = nq - 1 - (int)blockIdx.x, j_second = (int)blockIdx.x;
  const int n_items = (j_first == j_second) ? 1 : 2;

  if (warp == DKV_SM_WARPS && lane == 0) {
    for (int i = 0; i < 2; ++i) mbar_init(&kv_full[i], 1);
    for (int i = 0; i < DKV_STAGES; ++i) { mbar_init(&qdo_full[i], 33); mbar_init(&qdo_free[i], 1); }
    mbar_init(st_full, 1); mbar_init(st_free, DKV_SM_WARPS * 32); mbar_init(pds_full, DKV_SM_WARPS * 32); mbar_init(acc_done, 1);
    fence_barrier_init();
    for (int it = 0; it < n_items; ++it) {
      const int j = it ? j_second : j_first;
      mbar_expect_tx(&kv_full[it], 2 * TILE_BYTES);
      tma_load_3d(sKV + it * 2 * TILE_BYTES, &tmQKV, 0, k_slot, row_base + j * ATT_BK, &kv_full[it]);
      tma_load_3d(sKV + it * 2 * TILE_BYTES + TILE_BYTES, &tmQKV, 0, v_slot, row_base + j * ATT_BK, &kv_full[it]);
    }
  }
  if (warp == DKV_SM_WARPS + 1) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  const uint32_t tST = tmem, tdPT = tmem + 128, tPT = tmem + 256, tdST = tmem + 320, tdV = tmem + 384, tdK = tmem + 448;

  if (warp == DKV_SM_WARPS) {
    // ---------------- producer: Q / dO tiles by TMA, their lse / delta rows by the 32 lanes ----------------
    int n = 0;
    for (int item = 0; item < n_items; ++item) {
      const int j = item ? j_second : j_first;
      for (int hh = 0; hh < p.qpk; ++hh) {
        const int h = g * p.qpk + hh;
        for (int i = j; i < nq; ++i, ++n) {
          const int s = n % DKV_STAGES;
          // the statistics go to registers first (independent loads, in flight across the wait for the stage)
          const long long sbase = stat_index(p, b, h, row_base, 0);
          float lv[4], dv[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int tq = i * ATT_BQ + lane + 32 * k;
            const bool ok = tq < T_b;      // rows beyond the sequence (the next sequence's, or zero fill) contribute nothing
            lv[k] = ok ? __ldg(p.lse + sbase + tq) * LOG2E : INFINITY;   // +inf => p = 0 for padded query rows
            dv[k] = ok ? __ldg(p.delta + sbase + tq) : 0.f;
          }
          if (n >= DKV_STAGES) mbar_wait(&qdo_free[s], ((n / DKV_STAGES) - 1) & 1);
          if (lane == 0) {
            mbar_expect_tx(&qdo_full[s], 2 * TILE_BYTES);
            tma_load_3d(sQdO + s * 2 * TILE_BYTES, &tmQKV, 0, g * (p.qpk + 2) + hh, row_base + i * ATT_BQ, &qdo_full[s]);
            tma_load_3d(sQdO + s * 2 * TILE_BYTES + TILE_BYTES, &tmDO, 0, h, row_base + i * ATT_BQ, &qdo_full[s]);
          }
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            sStat[s * 256 + lane + 32 * k] = lv[k];
            sStat[s * 256 + 128 + lane + 32 * k] = dv[k];
          }
          mbar_arrive(&qdo_full[s]);
        }
      }
    }
  } else if (warp == DKV_SM_WARPS + 1) {
    // ---------------- MMA issuer: the whole warp walks the loop (uniform control flow and operands), one elected
    // lane issues -- a `lane == 0` region would make every descriptor a per-lane value and cost ~20 instructions per MMA
    int n = 0, nfree = 0, npds = 0;
    for (int item = 0; item < n_items; ++item) {
      const int j = item ? j_second : j_first;
      const int nit = p.qpk * (nq - j);
      const uint32_t k_addr = smem_u32(sKV + item * 2 * TILE_BYTES), v_addr = k_addr + TILE_BYTES;
      mbar_wait(&kv_full[item], 0);
      mbar_wait(&qdo_full[n % DKV_STAGES], (n / DKV_STAGES) & 1);
      if (n > 0) { mbar_wait(st_free, nfree & 1); ++nfree; }
      tc_fence_after();
      {
        const uint32_t q_addr = smem_u32(sQdO + (n % DKV_STAGES) * 2 * TILE_BYTES);
        if (elect_one()) {
          mma_qk(tST, k_addr, q_addr);
          mma_qk(tdPT, v_addr, q_addr + TILE_BYTES);
          tc_commit(st_full);
        }
        __syncwarp();
      }
      for (int it = 0; it < nit; ++it, ++n) {
        if (it + 1 < nit) {
          const int n1 = n + 1;
          TRC(20);
          mbar_wait(&qdo_full[n1 % DKV_STAGES], (n1 / DKV_STAGES) & 1);
          TRC(21);
          mbar_wait(st_free, nfree & 1); ++nfree;       // S^T / dP^T of iteration n are in registers
          TRC(22);
          tc_fence_after();
          const uint32_t q1_addr = smem_u32(sQdO + (n1 % DKV_STAGES) * 2 * TILE_BYTES);
          if (elect_one()) {
            mma_qk(tST, k_addr, q1_addr);
            mma_qk(tdPT, v_addr, q1_addr + TILE_BYTES);
            tc_commit(st_full);
          }
          __syncwarp();
        }
        TRC(23);
        mbar_wait(pds_full, npds & 1); ++npds;          // P^T / dS^T of iteration n are in TMEM
        TRC(24);
        tc_fence_after();
        const uint32_t q_addr = smem_u32(sQdO + (n % DKV_STAGES) * 2 * TILE_BYTES);
        const bool acc = it > 0;
        if (elect_one()) {
          mma_ts_k128(tdV, tPT, q_addr + TILE_BYTES, acc);
          mma_ts_k128(tdK, tdST, q_addr, acc);
          tc_commit(&qdo_free[n % DKV_STAGES]);
          tc_commit(acc_done);
        }
        __syncwarp();
        TRC(25);
      }
    }
  } else {
    // ---------------- softmax-backward warps: four threads per key row, 32 query columns each ----------------
    const int quad = warp & 3, qt = warp >> 2;
    const int r = quad * 32 + lane;
    const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
    const float sl2 = p.scale_log2;
    int n = 0, npds = 0, nacc = 0;
    for (int item = 0; item < n_items; ++item) {
      const int j = item ? j_second : j_first;
      const int n_qt = nq - j;
      const int nit = p.qpk * n_qt;
      for (int it = 0; it < nit; ++it, ++n) {
        const int s = n % DKV_STAGES;
        const int i = j + it % n_qt;
        mbar_wait(&qdo_full[s], (n / DKV_STAGES) & 1);      // lse / delta rows of this q tile are in smem
        mbar_wait(st_full, n & 1);
        TRC(1);
        tc_fence_after();
        uint32_t sv[32], dv[32];
        tmem_ld32(tST + lane_off + 32 * qt, sv);
        tmem_ld32(tdPT + lane_off + 32 * qt, dv);
        tmem_wait_ld();
        TRC(2);
        tc_fence_before();
        mbar_arrive(st_free);
        // diagonal tile (i == j): key row r sees query column c iff c >= r; per 32-column chunk qt vs quad
        const int mode = (i == j) ? (qt > quad ? 0 : (qt == quad ? 2 : 3)) : 0;
        const uint32_t stat_addr = smem_u32(sStat + s * 256 + 32 * qt);
        while (nacc < npds) { mbar_wait(acc_done, nacc & 1); ++nacc; }   // dV / dK MMAs of the previous iteration have read P^T / dS^T
        TRC(3);
        tc_fence_after();
#pragma unroll
        for (int hp = 0; hp < 2; ++hp) {
          uint32_t pp[8], ds[8];
          bwd_piece_mode<true, true>(mode, sv + 16 * hp, dv + 16 * hp, 16 * hp, lane, sl2, stat_addr + 64 * hp, 0.f, 0.f, pp, ds);
          tmem_st8(tPT + lane_off + 16 * qt + 8 * hp, pp);
          tmem_st8(tdST + lane_off + 16 * qt + 8 * hp, ds);
        }
        TRC(4);
        tmem_wait_st();
        TRC(5);
        tc_fence_before();
        mbar_arrive(pds_full); ++npds;
      }
      // ---- epilogue of this key tile: dV and dK (scaled, conjugate RoPE) rows to dqkv ----
      while (nacc < npds) { mbar_wait(acc_done, nacc & 1); ++nacc; }
      tc_fence_after();
      const int tk = j * ATT_BK + r;
      const bool valid = tk < T_b;
      bf16* rowp = p.dqkv + (long long)(row_base + (valid ? tk : 0)) * p.ld_qkv;
      uint32_t lo[8], hi[8];
      tmem_ld8(tdV + lane_off + 8 * qt, lo);
      tmem_ld8(tdV + lane_off + 32 + 8 * qt, hi);
      tmem_wait_ld();
      if (valid) store_rot_pairs(rowp + v_slot * HD, qt, lo, hi, 1.f, nullptr, nullptr, 0);
      tmem_ld8(tdK + lane_off + 8 * qt, lo);
      tmem_ld8(tdK + lane_off + 32 + 8 * qt, hi);
      tmem_wait_ld();
      tc_fence_before();
      if (valid) store_rot_pairs(rowp + k_slot * HD, qt, lo, hi, p.scale, p.rope_cos, p.rope_sin, tk);
    }
  }
  if (warp == 0 || warp >= DKV_SM_WARPS - 1) TRC_DUMP("dkv");
  tc_fence_before();
  __syncthreads();
  if (warp == DKV_SM_WARPS + 1) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

// =================================================================================================
// backward dQ.  CTA = (q tile pair {nq-1-x, x}, kv group, batch), one CTA per SM, flat loop over the qpk heads
// of the group and the key tiles j <= i of each query tile i (qpk*(nq+1) iterations per CTA, like dK/dV).
// The query row is the TMEM lane: S = Q K^T and dP = dO V^T (SS MMAs) -> sixteen softmax warps (four threads
// per row, 32 key columns each) -> dS = 2^(s - lse) (dP - delta) as bf16 back into TMEM -> dQ += dS K (TS MMA,
// K as the MN-major B operand).  K/V tiles run through a 3-deep ring, Q/dO tiles are double buffered over heads,
// and so is the dQ accumulator, so that the rows of head h (scaled, + the prefix part the prep kernel left in
// dqkv, conjugate RoPE) are written out while head h+1 is already being accumulated.
// =================================================================================================
constexpr int DQ_SM_WARPS = 16;
constexpr int DQ_THREADS = (DQ_SM_WARPS + 2) * 32;
constexpr int DQ_KV_STAGES = 3;
constexpr int DQ_OFF_KV = 4 * TILE_BYTES, DQ_OFF_BAR = (4 + 2 * DQ_KV_STAGES) * TILE_BYTES;
constexpr int DQ_SMEM = DQ_OFF_BAR + 256 + 1024;

__global__ void __launch_bounds__(DQ_THREADS, 1) attn_bwd_dq_kernel(const __grid_constant__ CUtensorMap tmQKV,
                                                                     const __grid_constant__ CUtensorMap tmDO, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQdO = smem;                     // [2 heads in flight][Q | dO]
  uint8_t* sKV = smem + DQ_OFF_KV;          // [DQ_KV_STAGES][K | V]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + DQ_OFF_BAR);
  uint64_t *qdo_full = bars /*[2]*/, *qdo_free = bars + 2 /*[2]*/, *kv_full = bars + 4 /*[STAGES]*/, *kv_free = bars + 4 + DQ_KV_STAGES,
           *sp_full = bars + 4 + 2 * DQ_KV_STAGES, *sp_free = sp_full + 1, *ds_full = sp_full + 2, *dsq_done = sp_full + 3,
           *dq_done = sp_full + 4 /*[2]*/;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sp_full + 6);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int g = blockIdx.y, b = blockIdx.z;
  const int row_base = p.cu_seqlens ? __ldg(p.cu_seqlens + b) : b * p.T;
  const int T_b = p.cu_seqlens ? __ldg(p.cu_seqlens + b + 1) - row_base : p.T;
  const int nq = (T_b + ATT_BQ - 1) / ATT_BQ;
  if ((int)blockIdx.x >= (nq + 1) / 2) return;      // the grid is sized for the longest sequence
  const int k_slot = g * (p.qpk + 2) + p.qpk, v_slot = k_slot + 1;
  const int i_first = nq - 1 - (int)blockIdx.x, i_second = (int)blockIdx.x;
  const int n_items = (i_first == i_second) ? 1 : 2;

  if (warp == DQ_SM_WARPS && lane == 0) {
    for (int i = 0; i < 2; ++i) { mbar_init(&qdo_full[i], 1); mbar_init(&qdo_free[i], 1); mbar_init(&dq_done[i], 1); }
    for (int i = 0; i < DQ_KV_STAGES; ++i) { mbar_init(&kv_full[i], 1); mbar_init(&kv_free[i], 1); }
    mbar_init(sp_full, 1); mbar_init(sp_free, DQ_SM_WARPS * 32); mbar_init(ds_full, DQ_SM_WARPS * 32); mbar_init(dsq_done, 1);
    fence_barrier_init();
  }
  if (warp == DQ_SM_WARPS + 1) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  const uint32_t tS = tmem, tdP = tmem + 128, tdS = tmem + 256, tdQ = tmem + 320 /* [2] x 64 */;

  if (warp == DQ_SM_WARPS) {
    // ---------------- TMA producer ----------------
    if (lane == 0) {
      int n = 0, hc = 0;
      for (int item = 0; item < n_items; ++item) {
        const int i = item ? i_second : i_first;
        for (int hh = 0; hh < p.qpk; ++hh, ++hc) {
          const int qs = hc & 1;
          if (hc >= 2) mbar_wait(&qdo_free[qs], ((hc >> 1) - 1) & 1);
          mbar_expect_tx(&qdo_full[qs], 2 * TILE_BYTES);
          tma_load_3d(sQdO + qs * 2 * TILE_BYTES, &tmQKV, 0, g * (p.qpk + 2) + hh, row_base + i * ATT_BQ, &qdo_full[qs]);
          tma_load_3d(sQdO + qs * 2 * TILE_BYTES + TILE_BYTES, &tmDO, 0, g * p.qpk + hh, row_base + i * ATT_BQ, &qdo_full[qs]);
          for (int j = 0; j <= i; ++j, ++n) {
            const int s = n % DQ_KV_STAGES;
            if (n >= DQ_KV_STAGES) mbar_wait(&kv_free[s], ((n / DQ_KV_STAGES) - 1) & 1);
            mbar_expect_tx(&kv_full[s], 2 * TILE_BYTES);
            tma_load_3d(sKV + s * 2 * TILE_BYTES, &tmQKV, 0, k_slot, row_base + j * ATT_BK, &kv_full[s]);
            tma_load_3d(sKV + s * 2 * TILE_BYTES + TILE_BYTES, &tmQKV, 0, v_slot, row_base + j * ATT_BK, &kv_full[s]);
          }
        }
      }
    }
  } else if (warp == DQ_SM_WARPS + 1) {
    // ---------------- MMA issuer (whole warp walks the loop, one elected lane issues) ----------------
    // iteration n = (item, head counter hc, key tile j); the S / dP MMAs of iteration n+1 go out as soon as the
    // softmax warps hold iteration n in registers, the dQ MMA of iteration n once its dS is in TMEM
    int n = 0, hc = 0;
    int item_nx = 0, hh_nx = 0, j_nx = 0, hc_nx = 0;          // the iteration whose S / dP is issued next
    auto issue_sdp = [&](int n_it) {
      const int i_nx = item_nx ? i_second : i_first;
      const int qs = hc_nx & 1, s = n_it % DQ_KV_STAGES;
      if (j_nx == 0) mbar_wait(&qdo_full[qs], (hc_nx >> 1) & 1);
      mbar_wait(&kv_full[s], (n_it / DQ_KV_STAGES) & 1);
      if (n_it > 0) mbar_wait(sp_free, (n_it - 1) & 1);      // S / dP of iteration n_it-1 are in registers
      tc_fence_after();
      const uint32_t q_addr = smem_u32(sQdO + qs * 2 * TILE_BYTES), kv_addr = smem_u32(sKV + s * 2 * TILE_BYTES);
      if (elect_one()) {
        mma_qk(tS, q_addr, kv_addr);
        mma_qk(tdP, q_addr + TILE_BYTES, kv_addr + TILE_BYTES);
        tc_commit(sp_full);
      }
      __syncwarp();
      // advance the look-ahead iterator
      if (++j_nx > i_nx) { j_nx = 0; ++hc_nx; if (++hh_nx == p.qpk) { hh_nx = 0; ++item_nx; } }
    };
    issue_sdp(0);
    for (int item = 0; item < n_items; ++item) {
      const int i = item ? i_second : i_first;
      for (int hh = 0; hh < p.qpk; ++hh, ++hc) {
        for (int j = 0; j <= i; ++j, ++n) {
          if (item_nx < n_items) issue_sdp(n + 1);
          mbar_wait(ds_full, n & 1);                          // dS of iteration n is in TMEM
          tc_fence_after();
          const int s = n % DQ_KV_STAGES;
          const uint32_t k_addr = smem_u32(sKV + s * 2 * TILE_BYTES);
          const bool acc = j > 0, last = (j == i);
          const uint32_t d_tmem = tdQ + 64 * (hc & 1);
          if (elect_one()) {
            mma_ts_k128(d_tmem, tdS, k_addr, acc);
            tc_commit(&kv_free[s]);
            tc_commit(dsq_done);
            if (last) { tc_commit(&dq_done[hc & 1]); tc_commit(&qdo_free[hc & 1]); }
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ---------------- softmax-backward warps: four threads per query row, 32 key columns each ----------------
    const int quad = warp & 3, qt = warp >> 2;
    const int r = quad * 32 + lane;
    const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
    const float sl2 = p.scale_log2;
    int n = 0, hc = 0;
    int nds = 0, ndsq = 0;          // ds_full arrivals made / dsq_done events waited for
    int pend_hc = -1, pend_tq = 0, pend_h = 0;   // head whose dQ rows still have to be written out

    auto write_dq = [&](int hcw, int tq, int h) {
      // rows of head h: dQ * scale (+ prefix part left in dqkv by the prep kernel), conjugate RoPE, bf16
      mbar_wait(&dq_done[hcw & 1], (hcw >> 1) & 1);
      tc_fence_after();
      uint32_t lo[8], hi[8];
      tmem_ld8(tdQ + 64 * (hcw & 1) + lane_off + 8 * qt, lo);
      tmem_ld8(tdQ + 64 * (hcw & 1) + lane_off + 32 + 8 * qt, hi);
      tmem_wait_ld();
      tc_fence_before();
      if (tq < T_b) {
        bf16* row = p.dqkv + (long long)(row_base + tq) * p.ld_qkv + (g * (p.qpk + 2) + (h - g * p.qpk)) * HD;
        if (p.prefix_len > 0) {
          float a[8], c[8];
          unpack8(*reinterpret_cast<const uint4*>(row + 8 * qt), a);
          unpack8(*reinterpret_cast<const uint4*>(row + 32 + 8 * qt), c);
          const float inv = 1.f / p.scale;
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            lo[k] = __float_as_uint(__uint_as_float(lo[k]) + a[k] * inv);
            hi[k] = __float_as_uint(__uint_as_float(hi[k]) + c[k] * inv);
          }
        }
        store_rot_pairs(row, qt, lo, hi, p.scale, p.rope_cos, p.rope_sin, tq);
      }
    };

    for (int item = 0; item < n_items; ++item) {
      const int i = item ? i_second : i_first;
      const int tq = i * ATT_BQ + r;
      const bool valid = tq < T_b;
      for (int hh = 0; hh < p.qpk; ++hh, ++hc) {
        const int h = g * p.qpk + hh;
        const long long stat_idx = stat_index(p, b, h, row_base, valid ? tq : 0);
        const float lse2 = valid ? __ldg(p.lse + stat_idx) * LOG2E : INFINITY;    // +inf => p = 0 on padded rows
        const float delta = valid ? __ldg(p.delta + stat_idx) : 0.f;
        for (int j = 0; j <= i; ++j, ++n) {
          mbar_wait(sp_full, n & 1);
          tc_fence_after();
          uint32_t sv[32], dv[32];
          tmem_ld32(tS + lane_off + 32 * qt, sv);
          tmem_ld32(tdP + lane_off + 32 * qt, dv);
          tmem_wait_ld();
          tc_fence_before();
          mbar_arrive(sp_free);
          // diagonal tile (j == i): query row r sees key column c iff c <= r; per 32-column chunk qt vs quad
          const int mode = (j == i) ? (qt < quad ? 0 : (qt == quad ? 1 : 3)) : 0;
          while (ndsq < nds) { mbar_wait(dsq_done, ndsq & 1); ++ndsq; }   // the dQ MMA of the previous iteration has read dS
          tc_fence_after();
#pragma unroll
          for (int hp = 0; hp < 2; ++hp) {
            uint32_t pp[8], ds[8];
            bwd_piece_mode<false, false>(mode, sv + 16 * hp, dv + 16 * hp, 16 * hp, lane, sl2, 0u, lse2, delta, pp, ds);
            tmem_st8(tdS + lane_off + 16 * qt + 8 * hp, ds);
          }
          tmem_wait_st();
          tc_fence_before();
          mbar_arrive(ds_full); ++nds;
          if (pend_hc >= 0 && pend_hc < hc) { write_dq(pend_hc, pend_tq, pend_h); pend_hc = -1; }
          if (j == i) {
            if (pend_hc >= 0) write_dq(pend_hc, pend_tq, pend_h);      // single-iteration heads: flush the older one first
            pend_hc = hc; pend_tq = tq; pend_h = h;
          }
        }
      }
    }
    if (pend_hc >= 0) write_dq(pend_hc, pend_tq, pend_h);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == DQ_SM_WARPS + 1) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

// =================================================================================================
// backward, step 0.  delta = rowsum(dO * y_causal) for the dQ and dK/dV kernels, and the whole backward of the
// gated prefix term (ger/adapter.py:129-146).  out = y + gate * ay  =>  delta = rowsum(dO*out) - gate*rowsum(dO*ay).
//
// attn_bwd_delta_kernel  (no prefix): pure streaming, 8 lanes per (row, head).
// attn_bwd_prefix_kernel (prefix):    CTA = (row tiles strided, head), tensor cores for every contraction:
//     S_a  = Q AK^T, dPV = dO AV^T          two N=16 MMAs -> TMEM; thread = row: softmax over the aT prefix keys,
//     delta, d_gate, dS_a = scale p gate (dPV - <p,dPV>), gP = gate p
//     dQ_a = dS_a AK                         dS_a as the TMEM A operand (K = 16) -> bf16 rows into dqkv (the dQ
//                                            kernel adds them before the RoPE rotation)
//     [dAK^T ; dAV^T] += [Q^T ; dO^T] [dS_a | gP]   one M=128 (two stacked MN-major tiles), N=32, K=128 MMA chain,
//                                            accumulated in TMEM over all row tiles of the CTA, then atomics.
// =================================================================================================
__global__ void __launch_bounds__(256) attn_bwd_delta_kernel(const AttnParams p) {
  const long long M = p.cu_seqlens ? p.total_rows : (long long)p.B * p.T;
  const int sub = threadIdx.x & 7;                      // 8 lanes x 16 bytes = one head of one row
  const long long pairs = M * p.H;
  for (long long pr = (long long)blockIdx.x * 32 + (threadIdx.x >> 3); pr < pairs; pr += (long long)gridDim.x * 32) {
    const long long m = pr / p.H;
    const int h = (int)(pr % p.H);
    float ov[8], dv[8];
    unpack8(*reinterpret_cast<const uint4*>(p.o + m * (p.H * HD) + h * HD + sub * 8), ov);
    unpack8(*reinterpret_cast<const uint4*>(p.dout + m * (p.H * HD) + h * HD + sub * 8), dv);
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc += ov[i] * dv[i];
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    acc += __shfl_xor_sync(0xffffffffu, acc, 4);
    if (sub == 0) p.delta[stat_index_flat(p, h, m)] = acc;
  }
}

constexpr int PFX_THREADS = 192;
constexpr int PFX_STAGE = 3 * TILE_BYTES;                 // Q | dO | O
constexpr int PFX_OFF_AK = 2 * PFX_STAGE, PFX_OFF_AV = PFX_OFF_AK + 2048, PFX_OFF_B2 = PFX_OFF_AV + 2048, PFX_OFF_BAR = PFX_OFF_B2 + 8192;
constexpr int PFX_SMEM = PFX_OFF_BAR + 256 + 1024;

__global__ void __launch_bounds__(PFX_THREADS, 2) attn_bwd_prefix_kernel(const __grid_constant__ CUtensorMap tmQKV,
                                                                          const __grid_constant__ CUtensorMap tmDO,
                                                                          const __grid_constant__ CUtensorMap tmO,
                                                                          const __grid_constant__ CUtensorMap tmPFX, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sAK = smem + PFX_OFF_AK;
  uint8_t* sAV = smem + PFX_OFF_AV;
  uint8_t* sB2 = smem + PFX_OFF_B2;                        // [32 n][128 rows] bf16, K-major: n < 16 dS_a, n >= 16 gP
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + PFX_OFF_BAR);
  uint64_t *full = bars /*[2]*/, *stage_free = bars + 2 /*[2]*/, *pfx_full = bars + 4, *sp_full = bars + 5, *dsp_full = bars + 6,
           *dq_full = bars + 7, *acc_done = bars + 8;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);
  __shared__ float s_gate_part[4];

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int h = blockIdx.y;
  const int g = h / p.qpk;
  const int q_slot = g * (p.qpk + 2) + (h % p.qpk), k_slot = g * (p.qpk + 2) + p.qpk, v_slot = k_slot + 1;
  const long long M = p.cu_seqlens ? p.total_rows : (long long)p.B * p.T;      // row-wise: sequence boundaries do not matter here
  const int ntiles = (int)((M + 127) / 128);
  const int my_tiles = ((int)blockIdx.x < ntiles) ? (ntiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

  if (warp == 4 && lane == 0) {
    for (int i = 0; i < 2; ++i) { mbar_init(&full[i], 1); mbar_init(&stage_free[i], 1); }
    mbar_init(pfx_full, 1); mbar_init(sp_full, 1); mbar_init(dsp_full, 128); mbar_init(dq_full, 1); mbar_init(acc_done, 1);
    fence_barrier_init();
    mbar_expect_tx(pfx_full, 4096);
    tma_load_3d(sAK, &tmPFX, 0, k_slot, 0, pfx_full);
    tma_load_3d(sAV, &tmPFX, 0, v_slot, 0, pfx_full);
  }
  if (warp == 5) { tmem_alloc(tmem_slot, 256); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  const uint32_t tSp = tmem, tdPV = tmem + 16, tdSp = tmem + 32, tdq = tmem + 64, tacc = tmem + 128;

  if (warp == 4) {
    // ---------------- TMA producer ----------------
    if (lane == 0) {
      for (int n = 0; n < my_tiles; ++n) {
        const int s = n & 1;
        const int row0 = ((int)blockIdx.x + n * (int)gridDim.x) * 128;
        if (n >= 2) mbar_wait(&stage_free[s], ((n >> 1) - 1) & 1);
        mbar_expect_tx(&full[s], 3 * TILE_BYTES);
        tma_load_3d(smem + s * PFX_STAGE, &tmQKV, 0, q_slot, row0, &full[s]);
        tma_load_3d(smem + s * PFX_STAGE + TILE_BYTES, &tmDO, 0, h, row0, &full[s]);
        tma_load_3d(smem + s * PFX_STAGE + 2 * TILE_BYTES, &tmO, 0, h, row0, &full[s]);
      }
    }
  } else if (warp == 5) {
    // ---------------- MMA issuer (uniform control flow, one elected lane issues) ----------------
    mbar_wait(pfx_full, 0);
    for (int n = 0; n < my_tiles; ++n) {
      const int s = n & 1;
      const uint32_t q_addr = smem_u32(smem + s * PFX_STAGE), do_addr = q_addr + TILE_BYTES;
      const uint32_t ak_addr = smem_u32(sAK), av_addr = smem_u32(sAV), b2_addr = smem_u32(sB2);
      mbar_wait(&full[s], (n >> 1) & 1);
      tc_fence_after();
      if (elect_one()) {
        constexpr uint32_t idesc16 = umma_idesc_bf16(128, 16, 0, 0);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          tc_mma_ss(tSp, umma_desc_sw128(q_addr + k * 32, 16, 1024), umma_desc_sw128(ak_addr + k * 32, 16, 1024), idesc16, k != 0);
          tc_mma_ss(tdPV, umma_desc_sw128(do_addr + k * 32, 16, 1024), umma_desc_sw128(av_addr + k * 32, 16, 1024), idesc16, k != 0);
        }
        tc_commit(sp_full);
      }
      __syncwarp();
      mbar_wait(dsp_full, n & 1);              // dS_a in TMEM, [dS_a | gP] in smem
      tc_fence_after();
      const bool acc = n > 0;
      if (elect_one()) {
        if (p.need_dqkv) {
          tc_mma_ts(tdq, tdSp, umma_desc_sw128(ak_addr, 8192, 1024), umma_idesc_bf16(128, 64, 0, 1), 0);
          tc_commit(dq_full);
        }
        constexpr uint32_t idesc_t = umma_idesc_bf16(128, 32, 1, 0);
#pragma unroll
        for (int k = 0; k < 8; ++k)     // A: 16 rows (K) x [64 d of Q | 64 d of dO] (M), MN-major; B: [32 n][16 rows], K-major
          tc_mma_ss(tacc, umma_desc_sw128(q_addr + k * 2048, TILE_BYTES, 1024),
                    umma_desc_sw128(b2_addr + (k >> 2) * 4096 + (k & 3) * 32, 16, 1024), idesc_t, acc || k != 0);
        tc_commit(&stage_free[s]);
        tc_commit(acc_done);
      }
      __syncwarp();
    }
  } else {
    // ---------------- row threads: thread = one token row of the 128-row tile ----------------
    const int r = warp * 32 + lane;
    const uint32_t lane_off = (uint32_t)(warp * 32) << 16;
    const int aT = p.prefix_len;
    const float gate = p.gate[h];
    float dgate_acc = 0.f;
    for (int n = 0; n < my_tiles; ++n) {
      const int s = n & 1;
      const long long m = (long long)((int)blockIdx.x + n * (int)gridDim.x) * 128 + r;
      const bool valid = m < M;
      mbar_wait(&full[s], (n >> 1) & 1);
      // rowsum(dO * out) from the staged tiles (128B-swizzled rows)
      float dot_out = 0.f;
      {
        const uint8_t* sdO = smem + s * PFX_STAGE + TILE_BYTES + r * 128;
        const uint8_t* sO = sdO + TILE_BYTES;
#pragma unroll
        for (int c8 = 0; c8 < 8; ++c8) {
          float ov[8], dv[8];
          unpack8(*reinterpret_cast<const uint4*>(sO + ((c8 ^ (r & 7)) << 4)), ov);
          unpack8(*reinterpret_cast<const uint4*>(sdO + ((c8 ^ (r & 7)) << 4)), dv);
#pragma unroll
          for (int i = 0; i < 8; ++i) dot_out += ov[i] * dv[i];
        }
      }
      mbar_wait(sp_full, n & 1);
      tc_fence_after();
      uint32_t sv[16], dv[16];
      tmem_ld16(tSp + lane_off, sv);
      tmem_ld16(tdPV + lane_off, dv);
      tmem_wait_ld();
      float mx = -INFINITY;
#pragma unroll
      for (int a = 0; a < MAX_PREFIX; ++a) if (a < aT) mx = fmaxf(mx, __uint_as_float(sv[a]));
      float pa[MAX_PREFIX], sum = 0.f;
#pragma unroll
      for (int a = 0; a < MAX_PREFIX; ++a) {
        pa[a] = (a < aT) ? fast_exp2((__uint_as_float(sv[a]) - mx) * p.scale_log2) : 0.f;
        sum += pa[a];
      }
      const float inv = 1.f / sum;
      float doay = 0.f;
#pragma unroll
      for (int a = 0; a < MAX_PREFIX; ++a) { pa[a] *= inv; doay += pa[a] * __uint_as_float(dv[a]); }
      if (valid) {
        p.delta[stat_index_flat(p, h, m)] = dot_out - gate * doay;
        dgate_acc += doay;
      }
      // dS_a (scaled) and gP, bf16: A operand of dQ_a (TMEM) and B operand of the dAK / dAV accumulation (smem)
      uint32_t dsp[8];
      float dsf[MAX_PREFIX], gpf[MAX_PREFIX];
#pragma unroll
      for (int a = 0; a < MAX_PREFIX; ++a) {
        dsf[a] = valid ? pa[a] * gate * (__uint_as_float(dv[a]) - doay) * p.scale : 0.f;
        gpf[a] = valid ? gate * pa[a] : 0.f;
      }
#pragma unroll
      for (int a = 0; a < 8; ++a) dsp[a] = pack_bf16x2(dsf[2 * a], dsf[2 * a + 1]);
      tmem_st8(tdSp + lane_off, dsp);
      if (n > 0) mbar_wait(acc_done, (n - 1) & 1);      // the previous tile's accumulation has read sB2
      {
        uint8_t* b2 = sB2 + (r >> 6) * 4096;
        const int col = r & 63;
#pragma unroll
        for (int a = 0; a < MAX_PREFIX; ++a) {
          *reinterpret_cast<bf16*>(b2 + sw128_offset(a, col)) = __float2bfloat16_rn(dsf[a]);
          *reinterpret_cast<bf16*>(b2 + sw128_offset(16 + a, col)) = __float2bfloat16_rn(gpf[a]);
        }
      }
      fence_proxy_async_smem();
      tmem_wait_st();
      tc_fence_before();
      mbar_arrive(dsp_full);
      if (p.need_dqkv) {
        mbar_wait(dq_full, n & 1);
        tc_fence_after();
        uint32_t q0[32], q1[32];
        tmem_ld32(tdq + lane_off, q0);
        tmem_ld32(tdq + lane_off + 32, q1);
        tmem_wait_ld();
        tc_fence_before();
        if (valid) {
          uint4* dq4 = reinterpret_cast<uint4*>(p.dqkv + m * p.ld_qkv + q_slot * HD);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float t8[8], u8[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) { t8[k] = __uint_as_float(q0[8 * i + k]); u8[k] = __uint_as_float(q1[8 * i + k]); }
            dq4[i] = pack8(t8);
            dq4[4 + i] = pack8(u8);
          }
        }
      }
    }
    // ---- d(prefix k / v) and d(gate): one atomic per element and CTA ----
    if (my_tiles > 0) {
      mbar_wait(acc_done, (my_tiles - 1) & 1);
      tc_fence_after();
      uint32_t acc[16];
      // lanes 0..63 hold dAK^T rows (d = r) in columns 0..15, lanes 64..127 dAV^T rows (d = r - 64) in columns 16..31
      tmem_ld16(tacc + lane_off + (r < 64 ? 0 : 16), acc);
      tmem_wait_ld();
      tc_fence_before();
      float* dst = p.d_prefix_kv + (long long)(r < 64 ? k_slot : v_slot) * HD + (r & 63);
#pragma unroll
      for (int a = 0; a < MAX_PREFIX; ++a)
        if (a < aT) atomicAdd(dst + (long long)a * p.ld_qkv, __uint_as_float(acc[a]));
    }
    dgate_acc = warp_sum(dgate_acc);
    if (lane == 0) s_gate_part[warp] = dgate_acc;
    named_bar_sync(1, 128);
    if (threadIdx.x == 0) atomicAdd(p.d_gate + h, s_gate_part[0] + s_gate_part[1] + s_gate_part[2] + s_gate_part[3]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) { tc_fence_after(); tmem_dealloc(tmem, 256); }
}

// =================================================================================================
// decode: one query per (sequence, head) against the KV cache -- HBM-bound (the cache is read once per token).
// Block = (sequence, kv group): the group's q heads (<= 8) share every K / V row, so a row is loaded ONCE and
// used by all of them (native GQA).  Thread = (dg, ks): dg = tid % 8 owns head dims [8 dg, 8 dg + 8) (one 16-byte
// load per row), ks = tid / 8 walks keys ks, ks + 32, ...  => a warp reads 4 whole 128-byte rows per load
// instruction, 4 loads in flight per thread.  Scores go through shared memory (fp32), the softmax is one warp per
// head, P V accumulates 8 heads x 8 dims per thread and is reduced over the 32 key slices at the end.
// =================================================================================================
constexpr int DEC_MAX_QPK = 8;
constexpr int DEC_THREADS = 256;
constexpr int DEC_FIXED_FLOATS = DEC_MAX_QPK * HD /*q*/ + 8 * 8 * HD /*per-warp partial outputs*/ + DEC_MAX_QPK * HD /*out*/ + 2 * DEC_MAX_QPK;

__global__ void __launch_bounds__(DEC_THREADS) attn_decode_kernel(const bf16* __restrict__ qkv, const bf16* __restrict__ ck,
                                                                  const bf16* __restrict__ cv, const int* __restrict__ lens, bf16* out,
                                                                  const bf16* __restrict__ prefix_kv, const float* __restrict__ gate,
                                                                  int max_len, int G, int qpk, int prefix_len, long long ld_qkv, float scale) {
  extern __shared__ float sdec[];
  float* sq = sdec;                                   // [8][64]   scaled queries
  float* spart = sq + DEC_MAX_QPK * HD;               // [8 warps][8 heads][64]
  float* sout = spart + 8 * 8 * HD;                   // [8][64]
  float* sinv = sout + DEC_MAX_QPK * HD;              // [8] 1 / sum, [8] spare
  float* sc = sinv + 2 * DEC_MAX_QPK;                 // [qpk][stride] scores / probabilities
  const int stride = max_len + 1;                     // odd stride: the 8 heads of one key land in different banks
  const int seq = blockIdx.x, g = blockIdx.y;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int dg = tid & 7, ks = tid >> 3;
  const int H = G * qpk;
  const int len = lens[seq];
  for (int i = tid; i < DEC_MAX_QPK * HD; i += DEC_THREADS) {
    const int h = i >> 6;
    sq[i] = h < qpk ? __bfloat162float(qkv[(long long)seq * ld_qkv + (g * (qpk + 2) + h) * HD + (i & 63)]) * scale : 0.f;
  }
  __syncthreads();
  const uint4* kbase = reinterpret_cast<const uint4*>(ck + ((long long)seq * G + g) * max_len * HD) + dg;   // row k: kbase[8 k]
  const uint4* vbase = reinterpret_cast<const uint4*>(cv + ((long long)seq * G + g) * max_len * HD) + dg;

  // ---- scores ----
  {
    float qr[DEC_MAX_QPK][8];
#pragma unroll
    for (int h = 0; h < DEC_MAX_QPK; ++h)
#pragma unroll
      for (int j = 0; j < 8; ++j) qr[h][j] = sq[h * HD + dg * 8 + j];
    for (int kb = 0; kb < len; kb += 128) {      // block-uniform trip count: the shuffles below need whole warps
      const int k0 = kb + ks;
      uint4 kv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int k = k0 + 32 * u;
        kv[u] = k < len ? __ldg(kbase + 8 * (long long)k) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int k = k0 + 32 * u;
        float kf[8];
        unpack8(kv[u], kf);
        float part[DEC_MAX_QPK];
#pragma unroll
        for (int h = 0; h < DEC_MAX_QPK; ++h) {
          float a = 0.f;
#pragma unroll
          for (int j = 0; j < 8; ++j) a = fmaf(kf[j], qr[h][j], a);
          part[h] = a;
        }
        // transposing butterfly over the 8 dim-group lanes of this key: 4 + 2 + 1 shuffles leave head dg's full
        // dot product in lane dg (each step halves the heads a lane still carries)
        const bool b2 = dg & 4, b1 = dg & 2, b0 = dg & 1;
        float k4[4], k2[2];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float recv = __shfl_xor_sync(0xffffffffu, b2 ? part[i] : part[i + 4], 4);
          k4[i] = (b2 ? part[i + 4] : part[i]) + recv;
        }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const float recv = __shfl_xor_sync(0xffffffffu, b1 ? k4[i] : k4[i + 2], 2);
          k2[i] = (b1 ? k4[i + 2] : k4[i]) + recv;
        }
        const float recv = __shfl_xor_sync(0xffffffffu, b0 ? k2[0] : k2[1], 1);
        const float mine = (b0 ? k2[1] : k2[0]) + recv;
        if (k < len && dg < qpk) sc[dg * stride + k] = mine;
      }
    }
  }
  __syncthreads();
  // ---- softmax: warp h owns head h ----
  if (warp < qpk) {
    float* row = sc + warp * stride;
    float mx = -INFINITY;
    for (int k = lane; k < len; k += 32) mx = fmaxf(mx, row[k]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int k = lane; k < len; k += 32) {
      const float e = __expf(row[k] - mx);
      row[k] = e;
      sum += e;
    }
    sum = warp_sum(sum);
    if (lane == 0) sinv[warp] = 1.f / sum;
  }
  __syncthreads();
  // ---- P V ----
  {
    float acc[DEC_MAX_QPK][8];
#pragma unroll
    for (int h = 0; h < DEC_MAX_QPK; ++h)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[h][j] = 0.f;
    for (int kb = 0; kb < len; kb += 128) {
      const int k0 = kb + ks;
      uint4 vv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int k = k0 + 32 * u;
        vv[u] = k < len ? __ldg(vbase + 8 * (long long)k) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int k = k0 + 32 * u;
        if (k < len) {
          float vf[8];
          unpack8(vv[u], vf);
#pragma unroll
          for (int h = 0; h < DEC_MAX_QPK; ++h) {
            const float pk = h < qpk ? sc[h * stride + k] : 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[h][j] = fmaf(pk, vf[j], acc[h][j]);
          }
        }
      }
    }
    // reduce over the 4 key slices of this warp (lane bits 3, 4), then over the 8 warps through shared memory
#pragma unroll
    for (int h = 0; h < DEC_MAX_QPK; ++h)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float a = acc[h][j];
        a += __shfl_xor_sync(0xffffffffu, a, 8);
        a += __shfl_xor_sync(0xffffffffu, a, 16);
        acc[h][j] = a;
      }
    if (lane < 8) {
#pragma unroll
      for (int h = 0; h < DEC_MAX_QPK; ++h)
#pragma unroll
        for (int j = 0; j < 8; ++j) spart[(warp * 8 + h) * HD + dg * 8 + j] = acc[h][j];
    }
  }
  __syncthreads();
  for (int i = tid; i < DEC_MAX_QPK * HD; i += DEC_THREADS) {
    float a = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) a += spart[w * 8 * HD + i];
    sout[i] = a * sinv[i >> 6];
  }
  __syncthreads();
  if (warp >= qpk) return;
  // ---- warp h finishes head h: + gate * softmax(q . prefix_k) prefix_v (its own softmax, ger/adapter.py:139-146) ----
  const int h = g * qpk + warp;
  float o0 = sout[warp * HD + lane], o1 = sout[warp * HD + lane + 32];
  if (prefix_len > 0) {
    const float q0 = sq[warp * HD + lane], q1 = sq[warp * HD + lane + 32];
    const long long koff = (long long)(g * (qpk + 2) + qpk) * HD;
    float ps[MAX_PREFIX];
    float pm = -INFINITY;
#pragma unroll
    for (int a = 0; a < MAX_PREFIX; ++a) {
      ps[a] = -INFINITY;
      if (a < prefix_len) {
        const bf16* kr = prefix_kv + (long long)a * ld_qkv + koff;
        float s = q0 * __bfloat162float(kr[lane]) + q1 * __bfloat162float(kr[lane + 32]);
        ps[a] = warp_sum(s);
        pm = fmaxf(pm, ps[a]);
      }
    }
    float psum = 0.f;
#pragma unroll
    for (int a = 0; a < MAX_PREFIX; ++a) {
      ps[a] = (a < prefix_len) ? __expf(ps[a] - pm) : 0.f;
      psum += ps[a];
    }
    const float gs = gate[h] / psum;
#pragma unroll
    for (int a = 0; a < MAX_PREFIX; ++a) {
      if (a < prefix_len) {
        const bf16* vr = prefix_kv + (long long)a * ld_qkv + koff + HD;
        o0 += gs * ps[a] * __bfloat162float(vr[lane]);
        o1 += gs * ps[a] * __bfloat162float(vr[lane + 32]);
      }
    }
  }
  bf16* orow = out + (long long)seq * H * HD + h * HD;
  orow[lane] = __float2bfloat16_rn(o0);
  orow[lane + 32] = __float2bfloat16_rn(o1);
}

// =================================================================================================
// decode on tensor cores.  The CUDA-core kernel above is ISSUE-bound (ncu: 9.3 M warp instructions per launch, 128 fp32
// FMAs per key and thread for the 8 heads that share a K / V row).  Here the 8 q heads of the group are the rows of an
// mma.sync m16n8k16 tile (rows 8..15 unused): scores S[head, key] = Q K^T with 8 keys per MMA column block, and
// O[head, dim] += P V with 16 keys per MMA k-step -- ~45x fewer issue slots, so the kernel is left with its K / V loads.
//   Q K^T : the k index of an MMA is only a summation index, so A and B use the same permuted k order and every
//           thread feeds the tensor core straight from one 16-byte load of its K row (dims 8q..8q+7 of a 32-dim block
//           supply k-slots {2q,2q+1 | 2q+8,2q+9} of two MMAs); no shared-memory staging for K.
//   P V   : B[k = key][n = dim] needs two KEYS of one dim per register: V rows go through a per-warp swizzled 2 KB
//           shared tile and come back transposed with ldmatrix.trans; P (fp32 in shared memory) is packed to bf16.
// Block = (sequence, kv group), 8 warps; warp w takes key blocks w, w + 8, ...
// =================================================================================================
// DM_NVB = V tiles (16 keys x 128 B, swizzled) per warp requested ahead by cp.async: 4 (all of V up to 512 keys in flight from
// the first instruction on) while two blocks still fit an SM, else 2 (long cache rows: the score buffer takes the room)
constexpr int dm_vt_bytes(int nvb) { return 8 * nvb * 2048; }
constexpr int dm_fixed_bytes(int nvb) {
  return dm_vt_bytes(nvb) /* also the per-warp partial outputs afterwards */ + (DEC_MAX_QPK * HD + 2 * DEC_MAX_QPK) * 4 + MAX_PREFIX * 256 /* prefix k | v */;
}

__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], uint32_t smem_addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_addr));
}

__device__ __forceinline__ void cp_async_16(uint32_t dst_smem, const void* src, int src_bytes) {   // src_bytes 0: zero-fill
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst_smem), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int DM_NVB>
__global__ void __launch_bounds__(DEC_THREADS, 2) attn_decode_mma_kernel(const bf16* __restrict__ qkv, const bf16* __restrict__ ck,
                                                                         const bf16* __restrict__ cv, const int* __restrict__ lens, bf16* out,
                                                                         const bf16* __restrict__ prefix_kv, const float* __restrict__ gate,
                                                                         int max_len, int G, int qpk, int prefix_len, long long ld_qkv, float scale) {
  extern __shared__ __align__(128) uint8_t sdm[];
  uint8_t* vt = sdm;                                                    // [8 warps][2][16 keys][128 B], later [8][8][64] floats
  float* spart = reinterpret_cast<float*>(sdm);
  float* sout = reinterpret_cast<float*>(sdm + dm_vt_bytes(DM_NVB));    // [8][64]
  float* sinv = sout + DEC_MAX_QPK * HD;                                // [8] 1 / sum, [8] spare
  bf16* spfx = reinterpret_cast<bf16*>(sinv + 2 * DEC_MAX_QPK);         // [MAX_PREFIX][k 64 | v 64] adapter prefix rows of this group
  float* sc = sinv + 2 * DEC_MAX_QPK + MAX_PREFIX * 64;                 // [qpk][stride] scores / probabilities
  const int stride = max_len + 20;                                      // multiple of 4 (8-byte P loads), rows 16 keys apart in banks
  const int seq = blockIdx.x, g = blockIdx.y;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int r = lane >> 2, q = lane & 3;                                // MMA fragment coordinates: row / column group, k pair
  const int H = G * qpk;
  pdl_launch_dependents();
  pdl_wait();
  const int len = lens[seq];
  const bf16* qrow = qkv + (long long)seq * ld_qkv + (long long)(g * (qpk + 2)) * HD;
  const bf16* kg = ck + ((long long)seq * G + g) * max_len * HD;
  const bf16* vg = cv + ((long long)seq * G + g) * max_len * HD;

  // the adapter prefix rows of this group (k | v, 256 B each) ride in the first cp.async group; they are used at the tail.
  // (Read straight from global memory they cost 2 x prefix_len dependent L2 round trips per warp: a quarter of the kernel.)
  if (tid < prefix_len * 16) {
    const int a = tid >> 4, c = tid & 15;
    cp_async_16(smem_u32(spfx + a * 128 + c * 8), prefix_kv + (long long)a * ld_qkv + (long long)(g * (qpk + 2) + qpk) * HD + c * 8, 16);
  }
  float q0 = 0.f, q1 = 0.f, gate_h = 0.f;
  if (warp < qpk && prefix_len > 0) {
    q0 = __bfloat162float(qrow[warp * HD + lane]) * scale;
    q1 = __bfloat162float(qrow[warp * HD + lane + 32]) * scale;
    gate_h = gate[g * qpk + warp];
  }
  // ---- V: this warp's first DM_NVB tiles (keys 16 (warp + 8 i) .. +15) are requested NOW, by cp.async straight into the
  // swizzled tiles, so that K and V share one memory round trip (up to 512 keys: all of V); rows at or beyond len are zero-filled
  const int nblk16 = (len + 15) >> 4;
  const uint32_t mytile_s = smem_u32(vt + warp * (DM_NVB * 2048));
  auto fetch_block = [&](int blk, int buf) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int e = lane + 32 * i, kin = e >> 3, c = e & 7, key = blk * 16 + kin;
      const bool ok = blk < nblk16 && key < len;
      cp_async_16(mytile_s + buf * 2048 + kin * 128 + ((c ^ (kin & 7)) << 4), reinterpret_cast<const uint4*>(vg + (long long)(ok ? key : 0) * HD) + c, ok ? 16 : 0);
    }
    cp_async_commit();
  };
#pragma unroll
  for (int i = 0; i < DM_NVB; ++i) fetch_block(warp + 8 * i, i);
  // ---- scores: A = Q (head r, unscaled bf16), B = K^T, fp32 accumulate, scale applied to the result ----
  {
    uint4 qa[2] = {make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};
    if (r < qpk) {
      qa[0] = *reinterpret_cast<const uint4*>(qrow + r * HD + 8 * q);
      qa[1] = *reinterpret_cast<const uint4*>(qrow + r * HD + 32 + 8 * q);
    }
    const int nblk = (len + 7) >> 3;
    for (int b0 = warp; b0 < nblk; b0 += 64) {        // 8 key blocks (16 independent 16-byte loads) in flight per thread
      uint4 kv[8][2];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int key = (b0 + 8 * u) * 8 + r;
        const bool ok = key < len;
        const uint4* kp = reinterpret_cast<const uint4*>(kg + (long long)key * HD + 8 * q);
        kv[u][0] = ok ? __ldg(kp) : make_uint4(0, 0, 0, 0);
        kv[u][1] = ok ? __ldg(kp + 4) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int key0 = (b0 + 8 * u) * 8;
        if (key0 < len) {                             // warp-uniform
          float c[4] = {0.f, 0.f, 0.f, 0.f};
          mma_bf16_16816(c, qa[0].x, 0u, qa[0].y, 0u, kv[u][0].x, kv[u][0].y);
          mma_bf16_16816(c, qa[0].z, 0u, qa[0].w, 0u, kv[u][0].z, kv[u][0].w);
          mma_bf16_16816(c, qa[1].x, 0u, qa[1].y, 0u, kv[u][1].x, kv[u][1].y);
          mma_bf16_16816(c, qa[1].z, 0u, qa[1].w, 0u, kv[u][1].z, kv[u][1].w);
          const int k = key0 + 2 * q;
          if (r < qpk) {
            if (k < len) sc[r * stride + k] = c[0] * scale;
            if (k + 1 < len) sc[r * stride + k + 1] = c[1] * scale;
          }
        }
      }
    }
  }
  __syncthreads();
  // ---- softmax: warp h owns head h; probabilities beyond len (up to the next multiple of 16) are zeroed for the P V tiles ----
  if (warp < qpk) {
    float* row = sc + warp * stride;
    float mx = -INFINITY;
    for (int k = lane; k < len; k += 32) mx = fmaxf(mx, row[k]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int k = lane; k < len; k += 32) {
      const float e = __expf(row[k] - mx);
      row[k] = e;
      sum += e;
    }
    sum = warp_sum(sum);
    if (lane == 0) sinv[warp] = 1.f / sum;
    if (lane < 16 && len + lane < ((len + 15) & ~15)) row[len + lane] = 0.f;
  }
  __syncthreads();
  // ---- P V: 16 keys per step; V rows -> swizzled per-warp tile -> ldmatrix.trans; P packed to bf16 from shared memory ----
  float acc[8][4];
#pragma unroll
  for (int nb = 0; nb < 8; ++nb)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[nb][j] = 0.f;
  {
    int it = 0;
    for (int blk = warp; blk < nblk16; blk += 8, ++it) {
      const int buf = it % DM_NVB;
      cp_async_wait<DM_NVB - 1>();                     // one group per tile, DM_NVB ahead: the oldest (this tile) has landed
      __syncwarp();
      // A = P[head r][16 keys]: a0 = keys 2q, 2q+1 ; a2 = keys 2q+8, 2q+9 (rows 8..15 of the tile are unused: a1 = a3 = 0)
      uint32_t a0 = 0u, a2 = 0u;
      if (r < qpk) {
        const float* pr = sc + r * stride + blk * 16 + 2 * q;
        const float2 p0 = *reinterpret_cast<const float2*>(pr), p1 = *reinterpret_cast<const float2*>(pr + 8);
        a0 = pack_bf16x2(p0.x, p0.y);
        a2 = pack_bf16x2(p1.x, p1.y);
      }
      const uint32_t tile_s = mytile_s + buf * 2048;
      // ldmatrix x4: lane supplies the row address of matrix m = lane / 8 (keys 8 (m & 1) .. +7, dim block 2 j + (m >> 1))
      const int m = lane >> 3, key = 8 * (m & 1) + (lane & 7);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int c = 2 * j + (m >> 1);
        uint32_t b[4];
        ldmatrix_x4_trans(b, tile_s + key * 128 + ((c ^ (key & 7)) << 4));
        mma_bf16_16816(acc[2 * j], a0, 0u, a2, 0u, b[0], b[1]);
        mma_bf16_16816(acc[2 * j + 1], a0, 0u, a2, 0u, b[2], b[3]);
      }
      __syncwarp();                                    // every lane has read the tile: it is refilled DM_NVB tiles ahead
      fetch_block(blk + 8 * DM_NVB, buf);
    }
    cp_async_wait<0>();                                // trailing (zero-fill) requests must land before the tiles become `spart`
  }
  __syncthreads();                                     // every warp is done with its V tiles: the region becomes the partial outputs
  if (r < DEC_MAX_QPK) {
#pragma unroll
    for (int nb = 0; nb < 8; ++nb)
      *reinterpret_cast<float2*>(spart + (warp * 8 + r) * HD + 8 * nb + 2 * q) = make_float2(acc[nb][0], acc[nb][1]);
  }
  __syncthreads();
  for (int i = tid; i < DEC_MAX_QPK * HD; i += DEC_THREADS) {
    float a = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) a += spart[w * 8 * HD + i];
    sout[i] = a * sinv[i >> 6];
  }
  __syncthreads();
  if (warp >= qpk) return;
  // ---- warp h finishes head h: + gate * softmax(q . prefix_k) prefix_v (its own softmax, ger/adapter.py:139-146) ----
  float pfx0 = 0.f, pfx1 = 0.f;
  if (prefix_len > 0) {
    float ps[MAX_PREFIX];
    float pm = -INFINITY;
#pragma unroll
    for (int a = 0; a < MAX_PREFIX; ++a) {
      ps[a] = -INFINITY;
      if (a < prefix_len) {
        ps[a] = warp_sum(q0 * __bfloat162float(spfx[a * 128 + lane]) + q1 * __bfloat162float(spfx[a * 128 + lane + 32]));
        pm = fmaxf(pm, ps[a]);
      }
    }
    float psum = 0.f;
#pragma unroll
    for (int a = 0; a < MAX_PREFIX; ++a) {
      ps[a] = (a < prefix_len) ? __expf(ps[a] - pm) : 0.f;
      psum += ps[a];
    }
    const float gs = gate_h / psum;
#pragma unroll
    for (int a = 0; a < MAX_PREFIX; ++a) {
      if (a < prefix_len) {
        pfx0 += gs * ps[a] * __bfloat162float(spfx[a * 128 + 64 + lane]);
        pfx1 += gs * ps[a] * __bfloat162float(spfx[a * 128 + 96 + lane]);
      }
    }
  }
  const int h = g * qpk + warp;
  const float o0 = sout[warp * HD + lane] + pfx0, o1 = sout[warp * HD + lane + 32] + pfx1;
  bf16* orow = out + (long long)seq * H * HD + h * HD;
  orow[lane] = __float2bfloat16_rn(o0);
  orow[lane + 32] = __float2bfloat16_rn(o1);
}

// ---------------------------------------------------------------------------------------------
// host
// ---------------------------------------------------------------------------------------------
static int make_head_map(CUtensorMap* map, const void* base, int slots, long long rows, long long ld, int box_rows = 128) {
  uint64_t dims[3] = {64, (uint64_t)slots, (uint64_t)rows};
  uint64_t strides[2] = {128, (uint64_t)ld * 2};
  uint32_t box[3] = {64, 1, (uint32_t)box_rows};
  return make_tensor_map(map, base, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

static void fill_common(AttnParams& p, int B, int T, int G, int qpk, int prefix_len) {
  p.B = B; p.T = T; p.G = G; p.qpk = qpk; p.H = G * qpk; p.slots = G * (qpk + 2);
  p.prefix_len = prefix_len;
  p.scale = 1.0f / sqrtf((float)HD);
  p.scale_log2 = p.scale * LOG2E;
  p.ld_qkv = (long long)p.slots * HD;
}

}  // namespace dh

using namespace dh;

extern "C" int dh_attention_fwd(const void* qkv, void* out, float* lse, const void* prefix_kv, const float* gate, int B, int T,
                                int n_groups, int q_per_kv, int prefix_len, void* stream) {
  DH_CHECK(qkv && out && lse && B > 0 && T > 0 && n_groups > 0 && q_per_kv > 0, "dh_attention_fwd: bad argument");
  DH_CHECK(prefix_len >= 0 && prefix_len <= MAX_PREFIX, "dh_attention_fwd: prefix_len %d > %d", prefix_len, MAX_PREFIX);
  DH_CHECK(prefix_len == 0 || (prefix_kv && gate), "dh_attention_fwd: prefix needs prefix_kv and gate");
  AttnParams p = {};
  fill_common(p, B, T, n_groups, q_per_kv, prefix_kv ? prefix_len : 0);
  p.qkv = (const bf16*)qkv; p.out = (bf16*)out; p.lse = lse;
  p.prefix_kv = (const bf16*)prefix_kv; p.gate = gate;
  CUtensorMap tm, tmp;
  if (make_head_map(&tm, qkv, p.slots, (long long)B * T, p.ld_qkv)) return -3;
  // prefix rows: a 16-row box over the prefix_len rows that exist, rows beyond are zero-filled by TMA
  if (make_head_map(&tmp, p.prefix_len > 0 ? prefix_kv : qkv, p.slots, p.prefix_len > 0 ? p.prefix_len : 1, p.ld_qkv, 16)) return -3;
  static bool attr = false;
  if (!attr) {
    DH_CUDA(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_SMEM));
    attr = true;
  }
  const int nq = (T + ATT_BQ - 1) / ATT_BQ;
  dim3 grid((nq + 1) / 2, p.H, B);
  attn_fwd_kernel<<<grid, FWD_THREADS, FWD_SMEM, (cudaStream_t)stream>>>(tm, tmp, p);
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_attention_fwd_varlen(const void* qkv, void* out, float* lse, const void* prefix_kv, const float* gate,
                                       const int32_t* cu_seqlens, int n_seq, int max_len, long long total_rows, int n_groups, int q_per_kv,
                                       int prefix_len, void* stream) {
  DH_CHECK(qkv && out && lse && cu_seqlens && n_seq > 0 && max_len > 0 && total_rows > 0 && n_groups > 0 && q_per_kv > 0,
           "dh_attention_fwd_varlen: bad argument");
  DH_CHECK(prefix_len >= 0 && prefix_len <= MAX_PREFIX, "dh_attention_fwd_varlen: prefix_len %d > %d", prefix_len, MAX_PREFIX);
  DH_CHECK(prefix_len == 0 || (prefix_kv && gate), "dh_attention_fwd_varlen: prefix needs prefix_kv and gate");
  AttnParams p = {};
  fill_common(p, n_seq, max_len, n_groups, q_per_kv, prefix_kv ? prefix_len : 0);
  p.qkv = (const bf16*)qkv; p.out = (bf16*)out; p.lse = lse;
  p.prefix_kv = (const bf16*)prefix_kv; p.gate = gate;
  p.cu_seqlens = cu_seqlens; p.total_rows = total_rows;
  CUtensorMap tm, tmp;
  if (make_head_map(&tm, qkv, p.slots, total_rows, p.ld_qkv)) return -3;
  if (make_head_map(&tmp, p.prefix_len > 0 ? prefix_kv : qkv, p.slots, p.prefix_len > 0 ? p.prefix_len : 1, p.ld_qkv, 16)) return -3;
  static bool attr = false;
  if (!attr) {
    DH_CUDA(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_SMEM));
    attr = true;
  }
  const int nq = (max_len + ATT_BQ - 1) / ATT_BQ;
  dim3 grid((nq + 1) / 2, p.H, n_seq);
  attn_fwd_kernel<<<grid, FWD_THREADS, FWD_SMEM, (cudaStream_t)stream>>>(tm, tmp, p);
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" size_t dh_attention_bwd_workspace(int B, int T, int n_groups, int q_per_kv) {
  return (size_t)B * T * n_groups * q_per_kv * sizeof(float);   // delta
}

// shared by the padded ([B, T]) and the packed (cu_seqlens) entry points: B = number of sequences, T = the longest one
static int attention_bwd_impl(const void* qkv, const void* out, const void* dout, const float* lse, const void* prefix_kv,
                              const float* gate, void* dqkv, float* d_prefix_kv, float* d_gate, int need_dqkv, const void* rope_cos,
                              const void* rope_sin, const int32_t* cu_seqlens, long long total_rows, int B, int T, int n_groups,
                              int q_per_kv, int prefix_len, void* workspace, size_t workspace_bytes, void* stream, const char* who) {
  DH_CHECK(qkv && out && dout && lse && B > 0 && T > 0 && total_rows > 0, "%s: bad argument", who);
  DH_CHECK(prefix_len >= 0 && prefix_len <= MAX_PREFIX, "%s: prefix_len %d > %d", who, prefix_len, MAX_PREFIX);
  DH_CHECK(workspace && workspace_bytes >= (size_t)total_rows * n_groups * q_per_kv * sizeof(float), "%s: workspace too small", who);
  DH_CHECK(!need_dqkv || dqkv, "%s: dqkv is NULL", who);
  const bool has_prefix = prefix_kv && prefix_len > 0;
  DH_CHECK(!has_prefix || (gate && d_prefix_kv && d_gate), "%s: prefix needs gate, d_prefix_kv, d_gate", who);
  DH_CHECK((rope_cos == nullptr) == (rope_sin == nullptr), "%s: pass both rope tables or neither", who);
  cudaStream_t s = (cudaStream_t)stream;
  AttnParams p = {};
  fill_common(p, B, T, n_groups, q_per_kv, has_prefix ? prefix_len : 0);
  p.qkv = (const bf16*)qkv; p.o = (const bf16*)out; p.dout = (const bf16*)dout; p.lse = const_cast<float*>(lse);
  p.prefix_kv = (const bf16*)prefix_kv; p.gate = gate;
  p.dqkv = (bf16*)dqkv; p.delta = (float*)workspace; p.d_prefix_kv = d_prefix_kv; p.d_gate = d_gate;
  p.need_dqkv = need_dqkv;
  p.rope_cos = (const bf16*)rope_cos; p.rope_sin = (const bf16*)rope_sin;
  p.cu_seqlens = cu_seqlens; p.total_rows = total_rows;
  const long long M = total_rows;
  if (!has_prefix) {
    const long long pairs = M * p.H;
    long long gx = (pairs + 31) / 32;
    if (gx > 148 * 16) gx = 148 * 16;
    attn_bwd_delta_kernel<<<(unsigned)gx, 256, 0, s>>>(p);
    DH_CUDA(cudaGetLastError());
  } else {
    CUtensorMap tQ, tDO, tO, tP;
    if (make_head_map(&tQ, qkv, p.slots, M, p.ld_qkv)) return -3;
    if (make_head_map(&tDO, dout, p.H, M, (long long)p.H * HD)) return -3;
    if (make_head_map(&tO, out, p.H, M, (long long)p.H * HD)) return -3;
    if (make_head_map(&tP, prefix_kv, p.slots, p.prefix_len, p.ld_qkv, 16)) return -3;
    static bool pattr = false;
    if (!pattr) {
      DH_CUDA(cudaFuncSetAttribute(attn_bwd_prefix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PFX_SMEM));
      pattr = true;
    }
    const int ntiles = (int)((M + 127) / 128);
    // two CTAs per SM and ONE wave: rounding the column count up (10 x 32 heads = 320 CTAs on 296 slots) left 24 CTAs to a second
    // wave that cost as much as the first
    int gx = (2 * sm_count()) / p.H;
    if (gx > ntiles) gx = ntiles;
    if (gx < 1) gx = 1;
    attn_bwd_prefix_kernel<<<dim3(gx, p.H), PFX_THREADS, PFX_SMEM, s>>>(tQ, tDO, tO, tP, p);
    DH_CUDA(cudaGetLastError());
  }
  if (!need_dqkv) return 0;
  CUtensorMap tmQKV, tmDO;
  if (make_head_map(&tmQKV, qkv, p.slots, M, p.ld_qkv)) return -3;
  if (make_head_map(&tmDO, dout, p.H, M, (long long)p.H * HD)) return -3;
  static bool attr = false;
  if (!attr) {
    DH_CUDA(cudaFuncSetAttribute(attn_bwd_dq_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DQ_SMEM));
    DH_CUDA(cudaFuncSetAttribute(attn_bwd_dkv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DKV_SMEM));
    attr = true;
  }
  const int nq = (T + ATT_BQ - 1) / ATT_BQ;
  attn_bwd_dkv_kernel<<<dim3((nq + 1) / 2, n_groups, B), DKV_THREADS, DKV_SMEM, s>>>(tmQKV, tmDO, p);
  DH_CUDA(cudaGetLastError());
  attn_bwd_dq_kernel<<<dim3((nq + 1) / 2, n_groups, B), DQ_THREADS, DQ_SMEM, s>>>(tmQKV, tmDO, p);
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, const void* prefix_kv,
                                const float* gate, void* dqkv, float* d_prefix_kv, float* d_gate, int need_dqkv,
                                const void* rope_cos, const void* rope_sin, int B, int T, int n_groups, int q_per_kv,
                                int prefix_len, void* workspace, size_t workspace_bytes, void* stream) {
  return attention_bwd_impl(qkv, out, dout, lse, prefix_kv, gate, dqkv, d_prefix_kv, d_gate, need_dqkv, rope_cos, rope_sin, nullptr,
                            (long long)B * T, B, T, n_groups, q_per_kv, prefix_len, workspace, workspace_bytes, stream, "dh_attention_bwd");
}

extern "C" int dh_attention_bwd_varlen(const void* qkv, const void* out, const void* dout, const float* lse, const void* prefix_kv,
                                       const float* gate, void* dqkv, float* d_prefix_kv, float* d_gate, int need_dqkv,
                                       const void* rope_cos, const void* rope_sin, const int32_t* cu_seqlens, int n_seq, int max_len,
                                       long long total_rows, int n_groups, int q_per_kv, int prefix_len, void* workspace,
                                       size_t workspace_bytes, void* stream) {
  DH_CHECK(cu_seqlens && n_seq > 0 && max_len > 0, "dh_attention_bwd_varlen: bad argument");
  return attention_bwd_impl(qkv, out, dout, lse, prefix_kv, gate, dqkv, d_prefix_kv, d_gate, need_dqkv, rope_cos, rope_sin, cu_seqlens,
                            total_rows, n_seq, max_len, n_groups, q_per_kv, prefix_len, workspace, workspace_bytes, stream,
                            "dh_attention_bwd_varlen");
}

extern "C" int dh_attention_decode_stream(const void* qkv, const void* cache_k, const void* cache_v, const int32_t* lens, void* out,
                                          const void* prefix_kv, const float* gate, int nseq, int max_len, int n_groups, int q_per_kv,
                                          int prefix_len, void* stream);

static int attention_decode_impl(int forced, const void* qkv, const void* cache_k, const void* cache_v, const int32_t* lens, void* out,
                                 const void* prefix_kv, const float* gate, int nseq, int max_len, int n_groups, int q_per_kv,
                                 int prefix_len, void* stream) {
  DH_CHECK(qkv && cache_k && cache_v && lens && out && nseq > 0, "dh_attention_decode: bad argument");
  DH_CHECK(q_per_kv <= DEC_MAX_QPK, "dh_attention_decode: q_per_kv %d > %d", q_per_kv, DEC_MAX_QPK);
  DH_CHECK(prefix_len >= 0 && prefix_len <= MAX_PREFIX, "dh_attention_decode: prefix_len too large");
  // default: the streamed kernel (csrc/decode_attn.cu: one persistent block of 15 flash-style warps + a TMA producer per SM, K / V as a
  // stream of 32-key chunks; measured 2.32 vs 2.63 ms per token step at 256 sequences, 1.56 vs 1.72 ms at 128); DH_DEC_ATTN=mma selects the whole-row mma.sync kernel
  // (also taken when the cache rows are not a multiple of 32 keys), DH_DEC_CUDA_CORES=1 the CUDA-core kernel (A/B runs)
  static int env_variant = -1;
  if (env_variant < 0) {
    const char* e = getenv("DH_DEC_CUDA_CORES");
    const char* a = getenv("DH_DEC_ATTN");
    env_variant = (e && e[0] == '1') ? 0 : (a && a[0] == 'm') ? 1 : 2;
    DH_CUDA(cudaFuncSetAttribute(attn_decode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    DH_CUDA(cudaFuncSetAttribute(attn_decode_mma_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    DH_CUDA(cudaFuncSetAttribute(attn_decode_mma_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  }
  const int variant = forced >= 0 ? forced : env_variant;
  if (variant == 2 && max_len % 32 == 0 && q_per_kv <= 8)
    return dh_attention_decode_stream(qkv, cache_k, cache_v, lens, out, prefix_kv, gate, nseq, max_len, n_groups, q_per_kv, prefix_len, stream);
  const long long ld = (long long)n_groups * (q_per_kv + 2) * HD;
  const bf16* pkv = (const bf16*)(prefix_len > 0 ? prefix_kv : nullptr);
  const int plen = prefix_kv ? prefix_len : 0;
  const float scale = 1.0f / sqrtf((float)HD);
  if (variant >= 1) {
    const size_t score_bytes = (size_t)q_per_kv * (max_len + 20) * sizeof(float);
    const bool deep = 2 * (dm_fixed_bytes(4) + score_bytes + 1024) <= 227 * 1024;      // two blocks per SM with 4 tiles per warp?
    const size_t smem = (size_t)dm_fixed_bytes(deep ? 4 : 2) + score_bytes;
    DH_CHECK(smem <= 200 * 1024, "dh_attention_decode: max_len %d too large for the score buffer", max_len);
    DH_CHECK((((uintptr_t)qkv | (uintptr_t)cache_k | (uintptr_t)cache_v | (uintptr_t)pkv) & 15) == 0, "dh_attention_decode: operands must be 16-byte aligned");
    DH_CHECK(plen <= MAX_PREFIX, "dh_attention_decode: prefix_len %d > %d", plen, MAX_PREFIX);
    // two blocks per SM (118 registers).  Three (an 80-register build, 220 B of spills) was measured SLOWER on 512 blocks:
    // 2.07 vs 1.89 ms per token step of 128 sequences.
    DH_CUDA(launch_pdl(deep ? attn_decode_mma_kernel<4> : attn_decode_mma_kernel<2>, dim3(nseq, n_groups), dim3(DEC_THREADS), smem,
                       (cudaStream_t)stream, (const bf16*)qkv, (const bf16*)cache_k, (const bf16*)cache_v, (const int*)lens, (bf16*)out, pkv,
                       (const float*)gate, max_len, n_groups, q_per_kv, plen, ld, scale));
  } else {
    const size_t smem = ((size_t)DEC_FIXED_FLOATS + (size_t)q_per_kv * (max_len + 1)) * sizeof(float);
    DH_CHECK(smem <= 200 * 1024, "dh_attention_decode: max_len %d too large for the score buffer", max_len);
    attn_decode_kernel<<<dim3(nseq, n_groups), DEC_THREADS, smem, (cudaStream_t)stream>>>(
        (const bf16*)qkv, (const bf16*)cache_k, (const bf16*)cache_v, lens, (bf16*)out, pkv, gate, max_len, n_groups, q_per_kv, plen, ld, scale);
  }
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_attention_decode(const void* qkv, const void* cache_k, const void* cache_v, const int32_t* lens, void* out,
                                   const void* prefix_kv, const float* gate, int nseq, int max_len, int n_groups, int q_per_kv,
                                   int prefix_len, void* stream) {
  return attention_decode_impl(-1, qkv, cache_k, cache_v, lens, out, prefix_kv, gate, nseq, max_len, n_groups, q_per_kv, prefix_len, stream);
}

// the whole-row mma.sync kernel, whatever DH_DEC_ATTN says (tests and A/B runs)
extern "C" int dh_attention_decode_rows(const void* qkv, const void* cache_k, const void* cache_v, const int32_t* lens, void* out,
                                        const void* prefix_kv, const float* gate, int nseq, int max_len, int n_groups, int q_per_kv,
                                        int prefix_len, void* stream) {
  return attention_decode_impl(1, qkv, cache_k, cache_v, lens, out, prefix_kv, gate, nseq, max_len, n_groups, q_per_kv, prefix_len, stream);
}
