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

#include "common.cuh"
#include "../../include/dualhyp.h"

namespace dh {

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

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 u;
  u.x = pack_bf16x2(f[0], f[1]); u.y = pack_bf16x2(f[2], f[3]); u.z = pack_bf16x2(f[4], f[5]); u.w = pack_bf16x2(f[6], f[7]);
  return u;
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
};

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
// forward
// =================================================================================================
constexpr int FWD_THREADS = 192;
constexpr int FWD_SMEM = 5 * TILE_BYTES + 2 * MAX_PREFIX * HD * 4 + 1024 + 128;

__global__ void __launch_bounds__(FWD_THREADS, 2) attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = smem + TILE_BYTES;
  uint8_t* sV = smem + 2 * TILE_BYTES;
  uint8_t* sP = smem + 3 * TILE_BYTES;   // two tiles
  float* sak = reinterpret_cast<float*>(smem + 5 * TILE_BYTES);
  float* sav = sak + MAX_PREFIX * HD;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sav + MAX_PREFIX * HD);
  uint64_t *q_full = bars, *k_full = bars + 1, *v_full = bars + 2, *s_full = bars + 3, *p_full = bars + 4, *o_done = bars + 5;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 6);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = (p.T + ATT_BQ - 1) / ATT_BQ;
  const int qi = nq - 1 - blockIdx.x;   // heavy tiles first
  const int h = blockIdx.y, b = blockIdx.z;
  const int g = h / p.qpk;
  const int q_slot = g * (p.qpk + 2) + (h % p.qpk), k_slot = g * (p.qpk + 2) + p.qpk, v_slot = k_slot + 1;
  const int row_base = b * p.T;
  const int q0 = qi * ATT_BQ;
  const int nkv = qi + 1;

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1); mbar_init(k_full, 1); mbar_init(v_full, 1);
    mbar_init(s_full, 1); mbar_init(p_full, 128); mbar_init(o_done, 1);
    fence_barrier_init();
    tma_prefetch_desc(&tmQKV);
  }
  if (warp == 5) { tmem_alloc(tmem_slot, 256); tmem_relinquish(); }
  if (p.prefix_len > 0) load_prefix_to_smem(p, g, sak, sav);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tO = tmem + 128;

  if (warp == 4) {
    if (lane == 0) {
      mbar_expect_tx(q_full, TILE_BYTES);
      tma_load_3d(sQ, &tmQKV, 0, q_slot, row_base + q0, q_full);
      for (int j = 0; j < nkv; ++j) {
        if (j > 0) mbar_wait(s_full, (j - 1) & 1);       // K(j-1) consumed by S(j-1)
        mbar_expect_tx(k_full, TILE_BYTES);
        tma_load_3d(sK, &tmQKV, 0, k_slot, row_base + j * ATT_BK, k_full);
        if (j > 0) mbar_wait(o_done, (j - 1) & 1);       // V(j-1) consumed by PV(j-1)
        mbar_expect_tx(v_full, TILE_BYTES);
        tma_load_3d(sV, &tmQKV, 0, v_slot, row_base + j * ATT_BK, v_full);
      }
    }
  } else if (warp == 5) {
    if (lane == 0) {
      mbar_wait(q_full, 0);
      for (int j = 0; j < nkv; ++j) {
        mbar_wait(k_full, j & 1);
        tc_fence_after();
        mma_qk(tS, smem_u32(sQ), smem_u32(sK));
        tc_commit(s_full);
        mbar_wait(p_full, j & 1);
        mbar_wait(v_full, j & 1);
        tc_fence_after();
        mma_pv(tO, smem_u32(sP), smem_u32(sV), j > 0);
        tc_commit(o_done);
      }
    }
  } else {
    // ---- softmax / epilogue: thread = query row ----
    const int r = warp * 32 + lane;
    const int tq = q0 + r;
    const uint32_t lane_off = (uint32_t)(warp * 32) << 16;
    float m2 = -INFINITY, l = 0.f;
    for (int j = 0; j < nkv; ++j) {
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      const int kv0 = j * ATT_BK;
      const bool diag = (j == qi);
      float mx = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        tmem_ld32(tS + lane_off + c * 32, v);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float s = __uint_as_float(v[i]);
          if (diag && kv0 + c * 32 + i > tq) s = -INFINITY;
          mx = fmaxf(mx, s);
        }
      }
      const float m_new = fmaxf(m2, mx * p.scale_log2);
      const float alpha = exp2f(m2 - m_new);
      if (j > 0) {
        mbar_wait(o_done, (j - 1) & 1);   // PV(j-1) done: O may be rescaled, P may be overwritten
        tc_fence_after();
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
          uint32_t v[32];
          tmem_ld32(tO + lane_off + c * 32, v);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) * alpha);
          tmem_st32(tO + lane_off + c * 32, v);
        }
        tmem_wait_st();
      }
      float rowsum = 0.f;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        tmem_ld32(tS + lane_off + c * 32, v);
        tmem_wait_ld();
        float f[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float e = exp2f(__uint_as_float(v[i]) * p.scale_log2 - m_new);
          if (diag && kv0 + c * 32 + i > tq) e = 0.f;
          f[i] = e;
          rowsum += e;
        }
        store_a_chunk(sP, r, c, f);
      }
      l = l * alpha + rowsum;
      m2 = m_new;
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(p_full);
    }
    mbar_wait(o_done, (nkv - 1) & 1);
    tc_fence_after();
    const float inv_l = 1.f / l;
    // gated prefix: its own softmax over the aT prefix keys (no RoPE on them), then gate[h] * (p_a @ av)
    float pw[MAX_PREFIX];
    const int aT = p.prefix_len;
    if (aT > 0) {
      float sp[MAX_PREFIX];
#pragma unroll
      for (int a = 0; a < MAX_PREFIX; ++a) sp[a] = 0.f;
#pragma unroll 1
      for (int c8 = 0; c8 < 8; ++c8) {
        float qv[8];
        unpack8(*reinterpret_cast<const uint4*>(sQ + r * 128 + ((c8 ^ (r & 7)) << 4)), qv);
#pragma unroll
        for (int a = 0; a < MAX_PREFIX; ++a) {
          if (a < aT) {
#pragma unroll
            for (int i = 0; i < 8; ++i) sp[a] += qv[i] * sak[a * HD + c8 * 8 + i];
          }
        }
      }
      float mxp = -INFINITY;
#pragma unroll
      for (int a = 0; a < MAX_PREFIX; ++a) if (a < aT) mxp = fmaxf(mxp, sp[a]);
      float sum = 0.f;
#pragma unroll
      for (int a = 0; a < MAX_PREFIX; ++a) {
        pw[a] = (a < aT) ? exp2f((sp[a] - mxp) * p.scale_log2) : 0.f;
        sum += pw[a];
      }
      const float gs = p.gate[h] / sum;
#pragma unroll
      for (int a = 0; a < MAX_PREFIX; ++a) pw[a] *= gs;
    }
    // TMEM loads are .sync.aligned: every lane issues them; only the global stores are predicated.
    const bool valid = tq < p.T;
    bf16* orow = p.out + (long long)(row_base + (valid ? tq : 0)) * (p.H * HD) + h * HD;
#pragma unroll 1
    for (int c = 0; c < 2; ++c) {
      uint32_t v[32];
      tmem_ld32(tO + lane_off + c * 32, v);
      tmem_wait_ld();
      float f[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]) * inv_l;
      if (aT > 0) {
#pragma unroll
        for (int a = 0; a < MAX_PREFIX; ++a) {
          if (a < aT) {
#pragma unroll
            for (int i = 0; i < 32; ++i) f[i] += pw[a] * sav[a * HD + c * 32 + i];
          }
        }
      }
      if (valid) {
        uint4* o4 = reinterpret_cast<uint4*>(orow + c * 32);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float t8[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) t8[k] = f[8 * i + k];
          o4[i] = pack8(t8);
        }
      }
    }
    if (valid) p.lse[((long long)b * p.H + h) * p.T + tq] = (m2 + log2f(l)) * LN2;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) { tc_fence_after(); tmem_dealloc(tmem, 256); }
}

// =================================================================================================
// backward, step 0: delta = rowsum(dO * y_causal) and the prefix / gate gradients.  CUDA cores.
// out = y + gate*ay  =>  rowsum(dO*y) = rowsum(dO*out) - gate*rowsum(dO*ay).
// block = 128 threads, thread = one token row of head blockIdx.y; blocks stride over row tiles.
// =================================================================================================
__global__ void __launch_bounds__(128) attn_bwd_prep_kernel(const AttnParams p) {
  __shared__ float sak[MAX_PREFIX * HD], sav[MAX_PREFIX * HD];
  __shared__ float sPA[4][32][MAX_PREFIX];   // gate * p[a]            (for d_av)
  __shared__ float sDS[4][32][MAX_PREFIX];   // scale * ds[a]          (for d_ak and dq)
  __shared__ float sgate[4];
  const int h = blockIdx.y;
  const int g = h / p.qpk;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int aT = p.prefix_len;
  const long long M = (long long)p.B * p.T;
  const int q_slot = g * (p.qpk + 2) + (h % p.qpk), k_slot = g * (p.qpk + 2) + p.qpk;
  if (aT > 0) load_prefix_to_smem(p, g, sak, sav);
  __syncthreads();
  const float gate = aT > 0 ? p.gate[h] : 0.f;
  float acc_v[MAX_PREFIX][2], acc_k[MAX_PREFIX][2];
#pragma unroll
  for (int a = 0; a < MAX_PREFIX; ++a) { acc_v[a][0] = acc_v[a][1] = acc_k[a][0] = acc_k[a][1] = 0.f; }
  float dgate_acc = 0.f;

  for (long long m0 = (long long)blockIdx.x * 128; m0 < M; m0 += (long long)gridDim.x * 128) {
    const long long m = m0 + threadIdx.x;
    const bool valid = m < M;
    float dot_out = 0.f;
    float sp[MAX_PREFIX], dpv[MAX_PREFIX];
#pragma unroll
    for (int a = 0; a < MAX_PREFIX; ++a) { sp[a] = 0.f; dpv[a] = 0.f; }
    if (valid) {
      const uint4* o4 = reinterpret_cast<const uint4*>(p.o + m * (p.H * HD) + h * HD);
      const uint4* d4 = reinterpret_cast<const uint4*>(p.dout + m * (p.H * HD) + h * HD);
      const uint4* q4 = reinterpret_cast<const uint4*>(p.qkv + m * p.ld_qkv + q_slot * HD);
#pragma unroll 1
      for (int c8 = 0; c8 < 8; ++c8) {
        float ov[8], dv[8];
        unpack8(o4[c8], ov);
        unpack8(d4[c8], dv);
#pragma unroll
        for (int i = 0; i < 8; ++i) dot_out += ov[i] * dv[i];
        if (aT > 0) {
          float qv[8];
          unpack8(q4[c8], qv);
#pragma unroll
          for (int a = 0; a < MAX_PREFIX; ++a) {
            if (a < aT) {
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                sp[a] += qv[i] * sak[a * HD + c8 * 8 + i];
                dpv[a] += dv[i] * sav[a * HD + c8 * 8 + i];
              }
            }
          }
        }
      }
    }
    float delta = dot_out;
    if (aT > 0) {
      float mx = -INFINITY, sum = 0.f, doay = 0.f;
#pragma unroll
      for (int a = 0; a < MAX_PREFIX; ++a) if (a < aT) mx = fmaxf(mx, sp[a]);
#pragma unroll
      for (int a = 0; a < MAX_PREFIX; ++a) {
        sp[a] = (a < aT) ? exp2f((sp[a] - mx) * p.scale_log2) : 0.f;
        sum += sp[a];
      }
      const float inv = 1.f / sum;
#pragma unroll
      for (int a = 0; a < MAX_PREFIX; ++a) { sp[a] *= inv; doay += sp[a] * dpv[a]; }
      delta = dot_out - gate * doay;
      if (valid) dgate_acc += doay;
      float ds[MAX_PREFIX];
#pragma unroll
      for (int a = 0; a < MAX_PREFIX; ++a) {
        ds[a] = valid ? sp[a] * gate * (dpv[a] - doay) * p.scale : 0.f;
        sPA[warp][lane][a] = valid ? gate * sp[a] : 0.f;
        sDS[warp][lane][a] = ds[a];
      }
      if (valid && p.need_dqkv) {   // dq_prefix = scale * sum_a ds[a] * ak[a]  (pre-rotation; the dQ kernel adds its part)
        uint4* dq4 = reinterpret_cast<uint4*>(p.dqkv + m * p.ld_qkv + q_slot * HD);
#pragma unroll 1
        for (int c8 = 0; c8 < 8; ++c8) {
          float f[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) f[i] = 0.f;
#pragma unroll
          for (int a = 0; a < MAX_PREFIX; ++a) {
            if (a < aT) {
#pragma unroll
              for (int i = 0; i < 8; ++i) f[i] += ds[a] * sak[a * HD + c8 * 8 + i];
            }
          }
          dq4[c8] = pack8(f);
        }
      }
      __syncwarp();
      // phase 2: lane owns head dims {lane, lane+32}; accumulate d_av += (gate p)^T dO, d_ak += (scale ds)^T q
      const long long mw = m0 + warp * 32;
#pragma unroll 1
      for (int rr = 0; rr < 32; ++rr) {
        const long long mr = mw + rr;
        if (mr >= M) break;
        const bf16* dr = p.dout + mr * (p.H * HD) + h * HD;
        const bf16* qr = p.qkv + mr * p.ld_qkv + q_slot * HD;
        const float d0 = __bfloat162float(dr[lane]), d1 = __bfloat162float(dr[lane + 32]);
        const float x0 = __bfloat162float(qr[lane]), x1 = __bfloat162float(qr[lane + 32]);
#pragma unroll
        for (int a = 0; a < MAX_PREFIX; ++a) {
          if (a < aT) {
            const float pa = sPA[warp][rr][a], da = sDS[warp][rr][a];
            acc_v[a][0] += pa * d0; acc_v[a][1] += pa * d1;
            acc_k[a][0] += da * x0; acc_k[a][1] += da * x1;
          }
        }
      }
      __syncwarp();
    }
    if (valid) {
      const int b = (int)(m / p.T), t = (int)(m % p.T);
      p.delta[((long long)b * p.H + h) * p.T + t] = delta;
    }
  }
  if (aT > 0) {
    float* dk = p.d_prefix_kv + (long long)k_slot * HD;
#pragma unroll
    for (int a = 0; a < MAX_PREFIX; ++a) {
      if (a < aT) {
        atomicAdd(dk + (long long)a * p.ld_qkv + lane, acc_k[a][0]);
        atomicAdd(dk + (long long)a * p.ld_qkv + lane + 32, acc_k[a][1]);
        atomicAdd(dk + (long long)a * p.ld_qkv + HD + lane, acc_v[a][0]);
        atomicAdd(dk + (long long)a * p.ld_qkv + HD + lane + 32, acc_v[a][1]);
      }
    }
    dgate_acc = warp_sum(dgate_acc);
    if (lane == 0) sgate[warp] = dgate_acc;
    __syncthreads();
    if (threadIdx.x == 0) atomicAdd(p.d_gate + h, sgate[0] + sgate[1] + sgate[2] + sgate[3]);
  }
}

// =================================================================================================
// backward dQ: CTA = (q tile, head, batch); loop over k tiles j <= qi.
//   S = Q K^T, dP = dO V^T  ->  dS = P * (dP - delta)  ->  dQ += dS K
// =================================================================================================
constexpr int BWD_THREADS = 192;
constexpr int DQ_SMEM = 6 * TILE_BYTES + 1024 + 128;

__global__ void __launch_bounds__(BWD_THREADS, 1) attn_bwd_dq_kernel(const __grid_constant__ CUtensorMap tmQKV,
                                                                      const __grid_constant__ CUtensorMap tmDO, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sdO = smem + TILE_BYTES;
  uint8_t* sK = smem + 2 * TILE_BYTES;
  uint8_t* sV = smem + 3 * TILE_BYTES;
  uint8_t* sdS = smem + 4 * TILE_BYTES;   // two tiles
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 6 * TILE_BYTES);
  uint64_t *q_full = bars, *k_full = bars + 1, *v_full = bars + 2, *sp_full = bars + 3, *ds_full = bars + 4, *dq_done = bars + 5;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 6);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = (p.T + ATT_BQ - 1) / ATT_BQ;
  const int qi = nq - 1 - blockIdx.x;
  const int h = blockIdx.y, b = blockIdx.z;
  const int g = h / p.qpk;
  const int q_slot = g * (p.qpk + 2) + (h % p.qpk), k_slot = g * (p.qpk + 2) + p.qpk, v_slot = k_slot + 1;
  const int row_base = b * p.T;
  const int q0 = qi * ATT_BQ;
  const int nkv = qi + 1;

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1); mbar_init(k_full, 1); mbar_init(v_full, 1);
    mbar_init(sp_full, 1); mbar_init(ds_full, 128); mbar_init(dq_done, 1);
    fence_barrier_init();
  }
  if (warp == 5) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tdP = tmem + 128, tdQ = tmem + 256;

  if (warp == 4) {
    if (lane == 0) {
      mbar_expect_tx(q_full, 2 * TILE_BYTES);
      tma_load_3d(sQ, &tmQKV, 0, q_slot, row_base + q0, q_full);
      tma_load_3d(sdO, &tmDO, 0, h, row_base + q0, q_full);
      for (int j = 0; j < nkv; ++j) {
        if (j > 0) mbar_wait(dq_done, (j - 1) & 1);      // K(j-1) consumed by dQ(j-1)
        mbar_expect_tx(k_full, TILE_BYTES);
        tma_load_3d(sK, &tmQKV, 0, k_slot, row_base + j * ATT_BK, k_full);
        if (j > 0) mbar_wait(sp_full, (j - 1) & 1);      // V(j-1) consumed by dP(j-1) (already true after dq_done)
        mbar_expect_tx(v_full, TILE_BYTES);
        tma_load_3d(sV, &tmQKV, 0, v_slot, row_base + j * ATT_BK, v_full);
      }
    }
  } else if (warp == 5) {
    if (lane == 0) {
      mbar_wait(q_full, 0);
      for (int j = 0; j < nkv; ++j) {
        mbar_wait(k_full, j & 1);
        mbar_wait(v_full, j & 1);
        tc_fence_after();
        mma_qk(tS, smem_u32(sQ), smem_u32(sK));
        mma_qk(tdP, smem_u32(sdO), smem_u32(sV));
        tc_commit(sp_full);
        mbar_wait(ds_full, j & 1);
        tc_fence_after();
        mma_pv(tdQ, smem_u32(sdS), smem_u32(sK), j > 0);
        tc_commit(dq_done);
      }
    }
  } else {
    const int r = warp * 32 + lane;
    const int tq = q0 + r;
    const bool valid = tq < p.T;
    const uint32_t lane_off = (uint32_t)(warp * 32) << 16;
    const long long stat_idx = ((long long)b * p.H + h) * p.T + (valid ? tq : 0);
    const float lse2 = p.lse[stat_idx] * LOG2E;
    const float delta = p.delta[stat_idx];
    for (int j = 0; j < nkv; ++j) {
      mbar_wait(sp_full, j & 1);
      tc_fence_after();
      if (j > 0) mbar_wait(dq_done, (j - 1) & 1);   // dS buffer free
      const int kv0 = j * ATT_BK;
      const bool diag = (j == qi);
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t s[32], d[32];
        tmem_ld32(tS + lane_off + c * 32, s);
        tmem_ld32(tdP + lane_off + c * 32, d);
        tmem_wait_ld();
        float f[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float pr = exp2f(__uint_as_float(s[i]) * p.scale_log2 - lse2);
          if ((diag && kv0 + c * 32 + i > tq) || !valid) pr = 0.f;
          f[i] = pr * (__uint_as_float(d[i]) - delta);
        }
        store_a_chunk(sdS, r, c, f);
      }
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(ds_full);
    }
    mbar_wait(dq_done, (nkv - 1) & 1);
    tc_fence_after();
    uint32_t v0[32], v1[32];
    tmem_ld32(tdQ + lane_off, v0);
    tmem_ld32(tdQ + lane_off + 32, v1);
    tmem_wait_ld();
    if (valid) {
      float lo[32], hi[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) { lo[i] = __uint_as_float(v0[i]) * p.scale; hi[i] = __uint_as_float(v1[i]) * p.scale; }
      bf16* dq = p.dqkv + (long long)(row_base + tq) * p.ld_qkv + q_slot * HD;
      if (p.prefix_len > 0) {   // prefix contribution written by the prep kernel
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float a[8], c[8];
          unpack8(reinterpret_cast<const uint4*>(dq)[i], a);
          unpack8(reinterpret_cast<const uint4*>(dq + 32)[i], c);
#pragma unroll
          for (int k = 0; k < 8; ++k) { lo[8 * i + k] += a[k]; hi[8 * i + k] += c[k]; }
        }
      }
      if (p.rope_cos) rope_conj_halves(lo, hi, p.rope_cos, p.rope_sin, tq);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float a[8], c[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) { a[k] = lo[8 * i + k]; c[k] = hi[8 * i + k]; }
        reinterpret_cast<uint4*>(dq)[i] = pack8(a);
        reinterpret_cast<uint4*>(dq + 32)[i] = pack8(c);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

// =================================================================================================
// backward dK/dV: CTA = (k tile j, kv head g, batch); loops over the q heads of the group and q tiles
// i >= j.  Thread = key row.  S^T = K Q^T, dP^T = V dO^T -> P^T, dS^T (smem) -> dV += P^T dO, dK += dS^T Q.
// Q/dO (+ their lse/delta) are double buffered.
// =================================================================================================
constexpr int DKV_SMEM = 10 * TILE_BYTES + 4 * 128 * 4 + 1024 + 128;

__global__ void __launch_bounds__(BWD_THREADS, 1) attn_bwd_dkv_kernel(const __grid_constant__ CUtensorMap tmQKV,
                                                                       const __grid_constant__ CUtensorMap tmDO, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem;
  uint8_t* sV = smem + TILE_BYTES;
  uint8_t* sQ = smem + 2 * TILE_BYTES;     // [2 stages]
  uint8_t* sdO = smem + 4 * TILE_BYTES;    // [2 stages]
  uint8_t* sPT = smem + 6 * TILE_BYTES;    // two tiles
  uint8_t* sdST = smem + 8 * TILE_BYTES;   // two tiles
  float* s_lse = reinterpret_cast<float*>(smem + 10 * TILE_BYTES);   // [2][128]
  float* s_delta = s_lse + 256;                                       // [2][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_delta + 256);
  uint64_t *kv_full = bars, *qdo_full = bars + 1 /*[2]*/, *qdo_free = bars + 3 /*[2]*/, *st_full = bars + 5, *pds_full = bars + 6,
           *acc_done = bars + 7;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = (p.T + ATT_BQ - 1) / ATT_BQ;
  const int j = blockIdx.x;               // k tile
  const int g = blockIdx.y, b = blockIdx.z;
  const int k_slot = g * (p.qpk + 2) + p.qpk, v_slot = k_slot + 1;
  const int row_base = b * p.T;
  const int kv0 = j * ATT_BK;
  const int n_qt = nq - j;                // q tiles j..nq-1
  const int nit = p.qpk * n_qt;

  if (threadIdx.x == 0) {
    mbar_init(kv_full, 1);
    for (int s = 0; s < 2; ++s) { mbar_init(&qdo_full[s], 1); mbar_init(&qdo_free[s], 1); }
    mbar_init(st_full, 1); mbar_init(pds_full, 128); mbar_init(acc_done, 1);
    fence_barrier_init();
  }
  if (warp == 5) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tST = tmem, tdPT = tmem + 128, tdV = tmem + 256, tdK = tmem + 320;

  if (warp == 4) {
    if (lane == 0) {
      mbar_expect_tx(kv_full, 2 * TILE_BYTES);
      tma_load_3d(sK, &tmQKV, 0, k_slot, row_base + kv0, kv_full);
      tma_load_3d(sV, &tmQKV, 0, v_slot, row_base + kv0, kv_full);
      for (int it = 0; it < nit; ++it) {
        const int s = it & 1;
        const int hh = it / n_qt, i = j + it % n_qt;
        const int h = g * p.qpk + hh;
        if (it >= 2) mbar_wait(&qdo_free[s], ((it >> 1) - 1) & 1);
        mbar_expect_tx(&qdo_full[s], 2 * TILE_BYTES);
        tma_load_3d(sQ + s * TILE_BYTES, &tmQKV, 0, g * (p.qpk + 2) + hh, row_base + i * ATT_BQ, &qdo_full[s]);
        tma_load_3d(sdO + s * TILE_BYTES, &tmDO, 0, h, row_base + i * ATT_BQ, &qdo_full[s]);
      }
    }
  } else if (warp == 5) {
    if (lane == 0) {
      mbar_wait(kv_full, 0);
      for (int it = 0; it < nit; ++it) {
        const int s = it & 1;
        mbar_wait(&qdo_full[s], (it >> 1) & 1);
        if (it > 0) mbar_wait(pds_full, (it - 1) & 1);   // S^T/dP^T of the previous iteration consumed
        tc_fence_after();
        mma_qk(tST, smem_u32(sK), smem_u32(sQ + s * TILE_BYTES));
        mma_qk(tdPT, smem_u32(sV), smem_u32(sdO + s * TILE_BYTES));
        tc_commit(st_full);
        mbar_wait(pds_full, it & 1);
        tc_fence_after();
        mma_pv(tdV, smem_u32(sPT), smem_u32(sdO + s * TILE_BYTES), it > 0);
        mma_pv(tdK, smem_u32(sdST), smem_u32(sQ + s * TILE_BYTES), it > 0);
        tc_commit(&qdo_free[s]);
        tc_commit(acc_done);
      }
    }
  } else {
    const int r = warp * 32 + lane;
    const int tk = kv0 + r;
    const uint32_t lane_off = (uint32_t)(warp * 32) << 16;
    for (int it = 0; it < nit; ++it) {
      const int s = it & 1;
      const int hh = it / n_qt, i = j + it % n_qt;
      const int h = g * p.qpk + hh;
      const int q0 = i * ATT_BQ;
      // stage lse / delta of this q tile (128 values each) for broadcast reads
      {
        const int tq = q0 + r;
        const bool ok = tq < p.T;
        const long long si = ((long long)b * p.H + h) * p.T + (ok ? tq : 0);
        s_lse[s * 128 + r] = ok ? p.lse[si] * LOG2E : INFINITY;   // +inf => p = 0 for padded query rows
        s_delta[s * 128 + r] = ok ? p.delta[si] : 0.f;
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      mbar_wait(st_full, it & 1);
      tc_fence_after();
      if (it > 0) mbar_wait(acc_done, (it - 1) & 1);   // P^T / dS^T buffers free
      const bool diag = (i == j);
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t sv[32], dv[32];
        tmem_ld32(tST + lane_off + c * 32, sv);
        tmem_ld32(tdPT + lane_off + c * 32, dv);
        tmem_wait_ld();
        float pt[32], ds[32];
#pragma unroll
        for (int ii = 0; ii < 32; ++ii) {
          const int col = c * 32 + ii;
          float pr = exp2f(__uint_as_float(sv[ii]) * p.scale_log2 - s_lse[s * 128 + col]);
          if (diag && tk > q0 + col) pr = 0.f;
          pt[ii] = pr;
          ds[ii] = pr * (__uint_as_float(dv[ii]) - s_delta[s * 128 + col]);
        }
        store_a_chunk(sPT, r, c, pt);
        store_a_chunk(sdST, r, c, ds);
      }
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(pds_full);
    }
    mbar_wait(acc_done, (nit - 1) & 1);
    tc_fence_after();
    uint32_t a0[32], a1[32];
    // dV
    tmem_ld32(tdV + lane_off, a0);
    tmem_ld32(tdV + lane_off + 32, a1);
    tmem_wait_ld();
    if (tk < p.T) {
      bf16* dv = p.dqkv + (long long)(row_base + tk) * p.ld_qkv + v_slot * HD;
#pragma unroll
      for (int i4 = 0; i4 < 4; ++i4) {
        float x[8], y[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) { x[k] = __uint_as_float(a0[8 * i4 + k]); y[k] = __uint_as_float(a1[8 * i4 + k]); }
        reinterpret_cast<uint4*>(dv)[i4] = pack8(x);
        reinterpret_cast<uint4*>(dv + 32)[i4] = pack8(y);
      }
    }
    // dK
    tmem_ld32(tdK + lane_off, a0);
    tmem_ld32(tdK + lane_off + 32, a1);
    tmem_wait_ld();
    if (tk < p.T) {
      float lo[32], hi[32];
#pragma unroll
      for (int i2 = 0; i2 < 32; ++i2) { lo[i2] = __uint_as_float(a0[i2]) * p.scale; hi[i2] = __uint_as_float(a1[i2]) * p.scale; }
      if (p.rope_cos) rope_conj_halves(lo, hi, p.rope_cos, p.rope_sin, tk);
      bf16* dk = p.dqkv + (long long)(row_base + tk) * p.ld_qkv + k_slot * HD;
#pragma unroll
      for (int i4 = 0; i4 < 4; ++i4) {
        float x[8], y[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) { x[k] = lo[8 * i4 + k]; y[k] = hi[8 * i4 + k]; }
        reinterpret_cast<uint4*>(dk)[i4] = pack8(x);
        reinterpret_cast<uint4*>(dk + 32)[i4] = pack8(y);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

// =================================================================================================
// decode: one query per (sequence, head) against the KV cache.  Block = (sequence, kv group); warp = one
// q head of the group; keys strided over lanes for the scores, head dims over lanes for P V.
// =================================================================================================
constexpr int DEC_MAX_QPK = 8;

__global__ void __launch_bounds__(256) attn_decode_kernel(const bf16* __restrict__ qkv, const bf16* __restrict__ ck,
                                                          const bf16* __restrict__ cv, const int* __restrict__ lens, bf16* out,
                                                          const bf16* __restrict__ prefix_kv, const float* __restrict__ gate, int max_len,
                                                          int G, int qpk, int prefix_len, long long ld_qkv, float scale) {
  extern __shared__ float sdec[];   // [qpk][max_len + MAX_PREFIX] scores
  const int seq = blockIdx.x, g = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp >= qpk) return;
  const int h = g * qpk + warp;
  const int H = G * qpk;
  const int len = lens[seq];
  float* sc = sdec + warp * (max_len + MAX_PREFIX);
  const bf16* q = qkv + (long long)seq * ld_qkv + (g * (qpk + 2) + warp) * HD;
  const float q0 = __bfloat162float(q[lane]) * scale, q1 = __bfloat162float(q[lane + 32]) * scale;
  const bf16* kbase = ck + ((long long)seq * G + g) * max_len * HD;
  const bf16* vbase = cv + ((long long)seq * G + g) * max_len * HD;
  // scores: each key needs a 64-dim dot; lanes hold dims {lane, lane+32}, reduce across the warp
  float mx = -INFINITY;
  for (int k = 0; k < len; ++k) {
    const bf16* kr = kbase + (long long)k * HD;
    float s = q0 * __bfloat162float(kr[lane]) + q1 * __bfloat162float(kr[lane + 32]);
    s = warp_sum(s);
    if (lane == 0) sc[k] = s;
    mx = fmaxf(mx, s);
  }
  __syncwarp();
  float sum = 0.f;
  for (int k = lane; k < len; k += 32) {
    const float e = __expf(sc[k] - mx);
    sc[k] = e;
    sum += e;
  }
  sum = warp_sum(sum);
  __syncwarp();
  float o0 = 0.f, o1 = 0.f;
  for (int k = 0; k < len; ++k) {
    const float pk = sc[k];
    const bf16* vr = vbase + (long long)k * HD;
    o0 += pk * __bfloat162float(vr[lane]);
    o1 += pk * __bfloat162float(vr[lane + 32]);
  }
  const float inv = 1.f / sum;
  o0 *= inv; o1 *= inv;
  if (prefix_len > 0) {
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

// ---------------------------------------------------------------------------------------------
// host
// ---------------------------------------------------------------------------------------------
static int make_head_map(CUtensorMap* map, const void* base, int slots, long long rows, long long ld) {
  uint64_t dims[3] = {64, (uint64_t)slots, (uint64_t)rows};
  uint64_t strides[2] = {128, (uint64_t)ld * 2};
  uint32_t box[3] = {64, 1, 128};
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
  CUtensorMap tm;
  if (make_head_map(&tm, qkv, p.slots, (long long)B * T, p.ld_qkv)) return -3;
  static bool attr = false;
  if (!attr) {
    DH_CUDA(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_SMEM));
    attr = true;
  }
  dim3 grid((T + ATT_BQ - 1) / ATT_BQ, p.H, B);
  attn_fwd_kernel<<<grid, FWD_THREADS, FWD_SMEM, (cudaStream_t)stream>>>(tm, p);
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" size_t dh_attention_bwd_workspace(int B, int T, int n_groups, int q_per_kv) {
  return (size_t)B * T * n_groups * q_per_kv * sizeof(float);   // delta
}

extern "C" int dh_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, const void* prefix_kv,
                                const float* gate, void* dqkv, float* d_prefix_kv, float* d_gate, int need_dqkv,
                                const void* rope_cos, const void* rope_sin, int B, int T, int n_groups, int q_per_kv,
                                int prefix_len, void* workspace, size_t workspace_bytes, void* stream) {
  DH_CHECK(qkv && out && dout && lse && B > 0 && T > 0, "dh_attention_bwd: bad argument");
  DH_CHECK(prefix_len >= 0 && prefix_len <= MAX_PREFIX, "dh_attention_bwd: prefix_len %d > %d", prefix_len, MAX_PREFIX);
  DH_CHECK(workspace && workspace_bytes >= dh_attention_bwd_workspace(B, T, n_groups, q_per_kv), "dh_attention_bwd: workspace too small");
  DH_CHECK(!need_dqkv || dqkv, "dh_attention_bwd: dqkv is NULL");
  const bool has_prefix = prefix_kv && prefix_len > 0;
  DH_CHECK(!has_prefix || (gate && d_prefix_kv && d_gate), "dh_attention_bwd: prefix needs gate, d_prefix_kv, d_gate");
  DH_CHECK((rope_cos == nullptr) == (rope_sin == nullptr), "dh_attention_bwd: pass both rope tables or neither");
  cudaStream_t s = (cudaStream_t)stream;
  AttnParams p = {};
  fill_common(p, B, T, n_groups, q_per_kv, has_prefix ? prefix_len : 0);
  p.qkv = (const bf16*)qkv; p.o = (const bf16*)out; p.dout = (const bf16*)dout; p.lse = const_cast<float*>(lse);
  p.prefix_kv = (const bf16*)prefix_kv; p.gate = gate;
  p.dqkv = (bf16*)dqkv; p.delta = (float*)workspace; p.d_prefix_kv = d_prefix_kv; p.d_gate = d_gate;
  p.need_dqkv = need_dqkv;
  p.rope_cos = (const bf16*)rope_cos; p.rope_sin = (const bf16*)rope_sin;
  const long long M = (long long)B * T;
  {
    int gx = (int)((M + 127) / 128);
    if (gx > 64) gx = 64;
    attn_bwd_prep_kernel<<<dim3(gx, p.H), 128, 0, s>>>(p);
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
  attn_bwd_dkv_kernel<<<dim3(nq, n_groups, B), BWD_THREADS, DKV_SMEM, s>>>(tmQKV, tmDO, p);
  DH_CUDA(cudaGetLastError());
  attn_bwd_dq_kernel<<<dim3(nq, p.H, B), BWD_THREADS, DQ_SMEM, s>>>(tmQKV, tmDO, p);
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_attention_decode(const void* qkv, const void* cache_k, const void* cache_v, const int32_t* lens, void* out,
                                   const void* prefix_kv, const float* gate, int nseq, int max_len, int n_groups, int q_per_kv,
                                   int prefix_len, void* stream) {
  DH_CHECK(qkv && cache_k && cache_v && lens && out && nseq > 0, "dh_attention_decode: bad argument");
  DH_CHECK(q_per_kv <= DEC_MAX_QPK, "dh_attention_decode: q_per_kv %d > %d", q_per_kv, DEC_MAX_QPK);
  DH_CHECK(prefix_len >= 0 && prefix_len <= MAX_PREFIX, "dh_attention_decode: prefix_len too large");
  const size_t smem = (size_t)q_per_kv * (max_len + MAX_PREFIX) * sizeof(float);
  DH_CHECK(smem <= 200 * 1024, "dh_attention_decode: max_len %d too large for the score buffer", max_len);
  static bool attr = false;
  if (!attr) {
    DH_CUDA(cudaFuncSetAttribute(attn_decode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr = true;
  }
  const long long ld = (long long)n_groups * (q_per_kv + 2) * HD;
  attn_decode_kernel<<<dim3(nseq, n_groups), 256, smem, (cudaStream_t)stream>>>(
      (const bf16*)qkv, (const bf16*)cache_k, (const bf16*)cache_v, lens, (bf16*)out, (const bf16*)(prefix_len > 0 ? prefix_kv : nullptr),
      gate, max_len, n_groups, q_per_kv, prefix_kv ? prefix_len : 0, ld, 1.0f / sqrtf((float)HD));
  DH_CUDA(cudaGetLastError());
  return 0;
}
