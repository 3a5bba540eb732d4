// One decode token step for all decoder layers as ONE persistent kernel (ger/model.py:163-268 per layer, driven by
// generate/base.py:57-80 one token at a time).  M <= 128 sequences, one new token each: the step is weight streaming
// (2.07 GB of bf16 weights + the K/V of the valid cache slots per token whatever the batch), so the design goal is that
// HBM never idles:
//
//   * one CTA per SM (148), launched once per token; the grid walks the layers' stages in lock step, separated by grid-wide
//     barriers (a monotonic arrive counter in global memory, acquire/release at gpu scope);
//   * a dedicated WEIGHT-PRODUCER warp streams this CTA's share of every GEMM's weight tiles through an 8-deep TMA ring and
//     never waits for a grid barrier: the weights are frozen, so while the other warps sit in a barrier, run a consumer stage
//     or the attention, the ring already holds the next GEMM's first tiles (19 MB in flight over the chip);
//   * every GEMM (QKV, attention proj, fc_1|fc_2, mlp proj) is cut into (128-column tile, 64-wide k-block) units and the
//     units are dealt out contiguously to the CTAs (stream-K): every SM streams the same number of weight bytes whatever
//     N / 128 is.  A CTA's run of units touches 1-3 output tiles; each (CTA, tile) segment leaves an fp32 partial tile
//     [128 x 128] in an L2-resident workspace; the tcgen05 accumulator is double buffered so the drain of one segment
//     overlaps the MMAs of the next;
//   * the stage after a GEMM sums a tile's segments in CTA order (deterministic) and applies what the reference does next:
//       QKV      -> RoPE + KV-cache write + the group's q row, inside the attention item that needs them (ger/model.py:216-259)
//       proj     -> residual add + RMSNorm (ger/model.py:185, ger/rmsnorm.py:17-21)
//       fc_1|fc_2-> silu(fc_1) * fc_2 (ger/model.py:313-316)
//       mlp proj -> residual add + the next layer's RMSNorm (or ln_f)
//   * attention: block = (sequence, kv group), the 8 q heads of the group are the rows of mma.sync m16n8k16 tiles over the
//     cached keys (native GQA), adapter prefix with its own softmax (ger/adapter.py:116-146) -- the algorithm of
//     attn_decode_mma_kernel (attention.cu) run as a stage of this kernel.
//
// Warp roles (352 threads): warps 0-7 compute (barriers, partial-tile drain, consumers, attention), warp 8 weight TMA,
// warp 9 activation TMA, warp 10 tcgen05.mma issue.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "../../include/dualhyp.h"

namespace dh {

constexpr int DK_CWARPS = 8;
constexpr int DK_CTHREADS = DK_CWARPS * 32;
constexpr int DK_THREADS = (DK_CWARPS + 3) * 32;
constexpr int DK_TILE_BYTES = 128 * 64 * 2;        // one [128 rows x 64 k] bf16 box
constexpr int DK_W_STAGES = 8;
constexpr int DK_A_STAGES = 2;
constexpr int DK_HD = 64;
constexpr int DK_MAX_QPK = 8;
constexpr int DK_MAX_PREFIX = 16;
constexpr int DK_MAX_LAYERS = 64;

struct DecodeLayer {
  CUtensorMap wqkv, wproj, w12, wmlp;   // [N, K] bf16 K-major, box {64 k, 128 n}
  const float* norm2;                   // this layer's norm_2 weight
  const float* norm_next;               // the next layer's norm_1 weight (ln_f after the last layer)
  bf16* cache_k;
  bf16* cache_v;
  const bf16* prefix_kv;                // adapter prefix K/V rows [prefix_len, qkv_out] (NULL: no adapter in this layer)
  const float* gate;
  long long pad_[2];
};
static_assert(sizeof(DecodeLayer) % 64 == 0, "tensor maps inside the layer array must stay 64-byte aligned");

struct DecodePlan {
  CUtensorMap a_n, a_y, a_act;          // activation operands [n_seq, K] bf16, box {64 k, 128 rows}
  int n_layer, n_seq, C, I, G, qpk, max_len, prefix_len;
  int maxseg[4];                        // partial-tile slices reserved per output tile, by GEMM (qkv, proj, fc12, mlp)
  int n_ctas, nvb;
  float eps, scale;
  float* x;                             // residual stream in / out, fp32 [n_seq, C]
  float* x2;
  bf16* n;                              // normalised activations in (layer 0's norm_1 applied) / scratch
  bf16* qkv;
  bf16* y;
  bf16* act;
  bf16* hidden;                         // ln_f output
  float* ws;                            // fp32 partial tiles
  const int* positions;                 // [n_seq] position of the token being decoded
  const bf16* rope_cos;
  const bf16* rope_sin;
  unsigned long long* bar;              // [0] arrive counter, [1] launch generation
  long long pad_[3];
};
static_assert(sizeof(DecodePlan) % 64 == 0, "the layer array follows the plan header and must stay 64-byte aligned");

// ---------------------------------------------------------------------------------------------
// small device helpers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void cbar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(DK_CTHREADS) : "memory"); }
__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ int ld_volatile_shared(const int* p) {
  int v;
  asm volatile("ld.volatile.shared.s32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
  return v;
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

static __device__ __noinline__ void dk_hang(const char* what, int a, int b) {
  printf("[dualhyp] decode step timed out in %s (block %d, %d / %d)\n", what, blockIdx.x, a, b);
  __trap();
}

// grid-wide barrier over the compute warps of all CTAs: every CTA arrives once on a monotonic counter
__device__ __forceinline__ void grid_barrier(unsigned long long* bar, unsigned long long target) {
  cbar_sync();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(bar, 1ULL);
    if (ld_acquire_u64(bar) < target) {
      const uint64_t t0 = globaltimer_ns();
      uint32_t spins = 0;
      while (ld_acquire_u64(bar) < target) {
        if (((++spins) & 0xFF) == 0 && globaltimer_ns() - t0 > DH_HANG_GUARD_NS) dk_hang("grid barrier", (int)(target & 0x7fffffff), 0);
      }
    }
    __threadfence();          // gpu-scope fence: also drops this SM's L1 lines (data other CTAs wrote since)
  }
  cbar_sync();
}

struct Phase {
  const CUtensorMap* w;
  const CUtensorMap* a;
  int tiles, nkb, maxseg, units;
};
__device__ __forceinline__ Phase get_phase(const DecodePlan* p, const DecodeLayer* L, int g) {
  Phase ph;
  if (g == 0) { ph.w = &L->wqkv; ph.a = &p->a_n; ph.tiles = p->G * (p->qpk + 2) * DK_HD / 128; ph.nkb = p->C / 64; }
  else if (g == 1) { ph.w = &L->wproj; ph.a = &p->a_y; ph.tiles = p->C / 128; ph.nkb = p->C / 64; }
  else if (g == 2) { ph.w = &L->w12; ph.a = &p->a_n; ph.tiles = 2 * p->I / 128; ph.nkb = p->C / 64; }
  else { ph.w = &L->wmlp; ph.a = &p->a_act; ph.tiles = p->C / 128; ph.nkb = p->I / 64; }
  ph.maxseg = p->maxseg[g];
  ph.units = ph.tiles * ph.nkb;
  return ph;
}
// contiguous run of (tile, k-block) units of CTA c: [unit_lo(c), unit_lo(c + 1))
__device__ __host__ __forceinline__ int unit_lo(int units, int c, int P) { return (int)((long long)units * c / P); }
// the CTA whose run holds unit u
__device__ __host__ __forceinline__ int unit_owner(int units, int u, int P) { return (int)((((long long)u + 1) * P - 1) / units); }

// ---------------------------------------------------------------------------------------------
// attention stage (one (sequence, kv group) item; 256 compute threads)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void dk_mma_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void dk_ldmatrix_x4_trans(uint32_t (&r)[4], uint32_t smem_addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_addr));
}
__device__ __forceinline__ void dk_cp_async_16(uint32_t dst_smem, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst_smem), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void dk_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void dk_cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__host__ __device__ constexpr int dk_vt_bytes(int nvb) { return 8 * nvb * 2048; }
__host__ __device__ constexpr int dk_att_fixed_bytes(int nvb) {
  return dk_vt_bytes(nvb) + (DK_MAX_QPK * DK_HD + 2 * DK_MAX_QPK) * 4 + DK_MAX_PREFIX * 256;
}

// Sum the QKV GEMM's partial tiles for row `seq`, columns of kv group g (qpk + 2 head slots), apply NeoX RoPE to the q and k
// slots at the row's position (the DH_EPI_QKV_ROPE math: bf16-rounded product, bf16 tables, pairs (j, j + 32)), write the
// bf16 qkv row segment and put the k / v slots into the KV cache at (seq, pos).  80 of the 256 threads work.
__device__ __forceinline__ void qkv_reduce_rope_cache(const DecodePlan* p, const DecodeLayer* L, const Phase& ph, int seq, int g, int pos, int tid) {
  const int slots = p->qpk + 2;
  const int N = p->G * slots * DK_HD;
  const int P = p->n_ctas;
  for (int t = tid; t < slots * 8; t += DK_CTHREADS) {
    const int slot = t >> 3, j = (t & 7) * 4;
    const int col0 = (g * slots + slot) * DK_HD + j;
    const int tile = col0 >> 7, within = col0 & 127;
    const int first = unit_owner(ph.units, tile * ph.nkb, P), last = unit_owner(ph.units, (tile + 1) * ph.nkb - 1, P);
    float4 lo = make_float4(0.f, 0.f, 0.f, 0.f), hi = lo;
    const float* base = p->ws + ((size_t)tile * ph.maxseg * 128 + seq) * 128 + within;
    for (int s = 0; s <= last - first; ++s) {
      const float4 a = __ldcg(reinterpret_cast<const float4*>(base + (size_t)s * 128 * 128));
      const float4 b = __ldcg(reinterpret_cast<const float4*>(base + (size_t)s * 128 * 128 + 32));
      lo.x += a.x; lo.y += a.y; lo.z += a.z; lo.w += a.w;
      hi.x += b.x; hi.y += b.y; hi.z += b.z; hi.w += b.w;
    }
    float l4[4] = {lo.x, lo.y, lo.z, lo.w}, h4[4] = {hi.x, hi.y, hi.z, hi.w};
    if (slot <= p->qpk) {
      const uint2 cu = *reinterpret_cast<const uint2*>(p->rope_cos + (long long)pos * 64 + j);   // table halves are identical (repeat(1, 2))
      const uint2 su = *reinterpret_cast<const uint2*>(p->rope_sin + (long long)pos * 64 + j);
      const float2 c01 = unpack_bf16x2(cu.x), c23 = unpack_bf16x2(cu.y), s01 = unpack_bf16x2(su.x), s23 = unpack_bf16x2(su.y);
      const float c[4] = {c01.x, c01.y, c23.x, c23.y}, sn[4] = {s01.x, s01.y, s23.x, s23.y};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float x1 = bf16_round(l4[i]), x2 = bf16_round(h4[i]);
        l4[i] = x1 * c[i] - x2 * sn[i];
        h4[i] = x2 * c[i] + x1 * sn[i];
      }
    }
    const uint2 lo2 = make_uint2(pack_bf16x2(l4[0], l4[1]), pack_bf16x2(l4[2], l4[3]));
    const uint2 hi2 = make_uint2(pack_bf16x2(h4[0], h4[1]), pack_bf16x2(h4[2], h4[3]));
    bf16* dst = p->qkv + (long long)seq * N + col0;
    *reinterpret_cast<uint2*>(dst) = lo2;
    *reinterpret_cast<uint2*>(dst + 32) = hi2;
    if (slot >= p->qpk) {
      bf16* cd = (slot == p->qpk ? L->cache_k : L->cache_v) + (((long long)seq * p->G + g) * p->max_len + pos) * DK_HD + j;
      *reinterpret_cast<uint2*>(cd) = lo2;
      *reinterpret_cast<uint2*>(cd + 32) = hi2;
    }
  }
}

template <int NVB>
__device__ __forceinline__ void attention_item(const DecodePlan* p, const DecodeLayer* L, uint8_t* sdm, int seq, int g, int len, int tid) {
  const int warp = tid >> 5, lane = tid & 31;
  const int qpk = p->qpk, G = p->G, max_len = p->max_len;
  const int prefix_len = L->prefix_kv ? p->prefix_len : 0;
  const float scale = p->scale;
  const long long ld_qkv = (long long)G * (qpk + 2) * DK_HD;
  uint8_t* vt = sdm;                                                    // [8 warps][NVB][16 keys][128 B], later [8][8][64] floats
  float* spart = reinterpret_cast<float*>(sdm);
  float* sout = reinterpret_cast<float*>(sdm + dk_vt_bytes(NVB));       // [8][64]
  float* sinv = sout + DK_MAX_QPK * DK_HD;                              // [8] 1 / sum, [8] spare
  bf16* spfx = reinterpret_cast<bf16*>(sinv + 2 * DK_MAX_QPK);          // [MAX_PREFIX][k 64 | v 64]
  float* sc = sinv + 2 * DK_MAX_QPK + DK_MAX_PREFIX * 64;               // [qpk][stride] scores / probabilities
  const int stride = max_len + 20;
  const int r = lane >> 2, q = lane & 3;
  const int H = G * qpk;
  const bf16* qrow = p->qkv + (long long)seq * ld_qkv + (long long)(g * (qpk + 2)) * DK_HD;
  const bf16* kg = L->cache_k + ((long long)seq * G + g) * max_len * DK_HD;
  const bf16* vg = L->cache_v + ((long long)seq * G + g) * max_len * DK_HD;

  if (tid < prefix_len * 16) {
    const int a = tid >> 4, c = tid & 15;
    dk_cp_async_16(smem_u32(spfx + a * 128 + c * 8), L->prefix_kv + (long long)a * ld_qkv + (long long)(g * (qpk + 2) + qpk) * DK_HD + c * 8, 16);
  }
  float q0 = 0.f, q1 = 0.f, gate_h = 0.f;
  if (warp < qpk && prefix_len > 0) {
    q0 = __bfloat162float(qrow[warp * DK_HD + lane]) * scale;
    q1 = __bfloat162float(qrow[warp * DK_HD + lane + 32]) * scale;
    gate_h = L->gate[g * qpk + warp];
  }
  const int nblk16 = (len + 15) >> 4;
  const uint32_t mytile_s = smem_u32(vt + warp * (NVB * 2048));
  auto fetch_block = [&](int blk, int buf) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int e = lane + 32 * i, kin = e >> 3, c = e & 7, key = blk * 16 + kin;
      const bool ok = blk < nblk16 && key < len;
      dk_cp_async_16(mytile_s + buf * 2048 + kin * 128 + ((c ^ (kin & 7)) << 4), reinterpret_cast<const uint4*>(vg + (long long)(ok ? key : 0) * DK_HD) + c, ok ? 16 : 0);
    }
    dk_cp_async_commit();
  };
#pragma unroll
  for (int i = 0; i < NVB; ++i) fetch_block(warp + 8 * i, i);
  // ---- scores ----
  {
    uint4 qa[2] = {make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};
    if (r < qpk) {
      qa[0] = __ldcg(reinterpret_cast<const uint4*>(qrow + r * DK_HD + 8 * q));
      qa[1] = __ldcg(reinterpret_cast<const uint4*>(qrow + r * DK_HD + 32 + 8 * q));
    }
    const int nblk = (len + 7) >> 3;
    for (int b0 = warp; b0 < nblk; b0 += 64) {
      uint4 kv[8][2];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int key = (b0 + 8 * u) * 8 + r;
        const bool ok = key < len;
        const uint4* kp = reinterpret_cast<const uint4*>(kg + (long long)key * DK_HD + 8 * q);
        kv[u][0] = ok ? __ldcg(kp) : make_uint4(0, 0, 0, 0);
        kv[u][1] = ok ? __ldcg(kp + 4) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int key0 = (b0 + 8 * u) * 8;
        if (key0 < len) {
          float c[4] = {0.f, 0.f, 0.f, 0.f};
          dk_mma_16816(c, qa[0].x, 0u, qa[0].y, 0u, kv[u][0].x, kv[u][0].y);
          dk_mma_16816(c, qa[0].z, 0u, qa[0].w, 0u, kv[u][0].z, kv[u][0].w);
          dk_mma_16816(c, qa[1].x, 0u, qa[1].y, 0u, kv[u][1].x, kv[u][1].y);
          dk_mma_16816(c, qa[1].z, 0u, qa[1].w, 0u, kv[u][1].z, kv[u][1].w);
          const int k = key0 + 2 * q;
          if (r < qpk) {
            if (k < len) sc[r * stride + k] = c[0] * scale;
            if (k + 1 < len) sc[r * stride + k + 1] = c[1] * scale;
          }
        }
      }
    }
  }
  cbar_sync();
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
  cbar_sync();
  float acc[8][4];
#pragma unroll
  for (int nb = 0; nb < 8; ++nb)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[nb][j] = 0.f;
  {
    int it = 0;
    for (int blk = warp; blk < nblk16; blk += 8, ++it) {
      const int buf = it % NVB;
      dk_cp_async_wait<NVB - 1>();
      __syncwarp();
      uint32_t a0 = 0u, a2 = 0u;
      if (r < qpk) {
        const float* pr = sc + r * stride + blk * 16 + 2 * q;
        const float2 p0 = *reinterpret_cast<const float2*>(pr), p1 = *reinterpret_cast<const float2*>(pr + 8);
        a0 = pack_bf16x2(p0.x, p0.y);
        a2 = pack_bf16x2(p1.x, p1.y);
      }
      const uint32_t tile_s = mytile_s + buf * 2048;
      const int m = lane >> 3, key = 8 * (m & 1) + (lane & 7);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int c = 2 * j + (m >> 1);
        uint32_t b[4];
        dk_ldmatrix_x4_trans(b, tile_s + key * 128 + ((c ^ (key & 7)) << 4));
        dk_mma_16816(acc[2 * j], a0, 0u, a2, 0u, b[0], b[1]);
        dk_mma_16816(acc[2 * j + 1], a0, 0u, a2, 0u, b[2], b[3]);
      }
      __syncwarp();
      fetch_block(blk + 8 * NVB, buf);
    }
    dk_cp_async_wait<0>();
  }
  cbar_sync();
  if (r < DK_MAX_QPK) {
#pragma unroll
    for (int nb = 0; nb < 8; ++nb)
      *reinterpret_cast<float2*>(spart + (warp * 8 + r) * DK_HD + 8 * nb + 2 * q) = make_float2(acc[nb][0], acc[nb][1]);
  }
  cbar_sync();
  for (int i = tid; i < DK_MAX_QPK * DK_HD; i += DK_CTHREADS) {
    float a = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) a += spart[w * 8 * DK_HD + i];
    sout[i] = a * sinv[i >> 6];
  }
  cbar_sync();
  if (warp < qpk) {
    float pfx0 = 0.f, pfx1 = 0.f;
    if (prefix_len > 0) {
      float ps[DK_MAX_PREFIX];
      float pm = -INFINITY;
#pragma unroll
      for (int a = 0; a < DK_MAX_PREFIX; ++a) {
        ps[a] = -INFINITY;
        if (a < prefix_len) {
          ps[a] = warp_sum(q0 * __bfloat162float(spfx[a * 128 + lane]) + q1 * __bfloat162float(spfx[a * 128 + lane + 32]));
          pm = fmaxf(pm, ps[a]);
        }
      }
      float psum = 0.f;
#pragma unroll
      for (int a = 0; a < DK_MAX_PREFIX; ++a) {
        ps[a] = (a < prefix_len) ? __expf(ps[a] - pm) : 0.f;
        psum += ps[a];
      }
      const float gs = gate_h / psum;
#pragma unroll
      for (int a = 0; a < DK_MAX_PREFIX; ++a) {
        if (a < prefix_len) {
          pfx0 += gs * ps[a] * __bfloat162float(spfx[a * 128 + 64 + lane]);
          pfx1 += gs * ps[a] * __bfloat162float(spfx[a * 128 + 96 + lane]);
        }
      }
    }
    const int h = g * qpk + warp;
    const float o0 = sout[warp * DK_HD + lane] + pfx0, o1 = sout[warp * DK_HD + lane + 32] + pfx1;
    bf16* orow = p->y + (long long)seq * H * DK_HD + h * DK_HD;
    orow[lane] = __float2bfloat16_rn(o0);
    orow[lane + 32] = __float2bfloat16_rn(o1);
  }
  cbar_sync();      // the shared-memory region is re-used by the next item
}

// ---------------------------------------------------------------------------------------------
// consumer stages
// ---------------------------------------------------------------------------------------------
// row `row`: x_out = resid + bf16(sum of the GEMM's partial tiles), y = rmsnorm(x_out) * w  (256 threads, C <= 8192)
__device__ __forceinline__ void resid_norm_row(const DecodePlan* p, const Phase& ph, int row, const float* resid, float* x_out, bf16* y,
                                               const float* w, float* red, int tid) {
  const int C = p->C, P = p->n_ctas;
  float v[4][8];
  float ss = 0.f;
#pragma unroll
  for (int rep = 0; rep < 4; ++rep) {
    const int c0 = rep * 2048 + tid * 8;
    if (c0 < C) {
      const int tile = c0 >> 7, within = c0 & 127;
      const int first = unit_owner(ph.units, tile * ph.nkb, P), last = unit_owner(ph.units, (tile + 1) * ph.nkb - 1, P);
      const float* base = p->ws + ((size_t)tile * ph.maxseg * 128 + row) * 128 + within;
      float acc[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] = 0.f;
      for (int s = 0; s <= last - first; ++s) {
        const float4 a = __ldcg(reinterpret_cast<const float4*>(base + (size_t)s * 128 * 128));
        const float4 b = __ldcg(reinterpret_cast<const float4*>(base + (size_t)s * 128 * 128 + 4));
        acc[0] += a.x; acc[1] += a.y; acc[2] += a.z; acc[3] += a.w; acc[4] += b.x; acc[5] += b.y; acc[6] += b.z; acc[7] += b.w;
      }
      const float4 r0 = __ldcg(reinterpret_cast<const float4*>(resid + (size_t)row * C + c0));
      const float4 r1 = __ldcg(reinterpret_cast<const float4*>(resid + (size_t)row * C + c0 + 4));
      const float rr[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        v[rep][i] = rr[i] + bf16_round(acc[i]);
        ss += v[rep][i] * v[rep][i];
      }
      *reinterpret_cast<float4*>(x_out + (size_t)row * C + c0) = make_float4(v[rep][0], v[rep][1], v[rep][2], v[rep][3]);
      *reinterpret_cast<float4*>(x_out + (size_t)row * C + c0 + 4) = make_float4(v[rep][4], v[rep][5], v[rep][6], v[rep][7]);
    }
  }
  ss = warp_sum(ss);
  cbar_sync();
  if ((tid & 31) == 0) red[tid >> 5] = ss;
  cbar_sync();
  float tot = 0.f;
#pragma unroll
  for (int i = 0; i < DK_CWARPS; ++i) tot += red[i];
  const float rs = rsqrtf(tot / (float)C + p->eps);
#pragma unroll
  for (int rep = 0; rep < 4; ++rep) {
    const int c0 = rep * 2048 + tid * 8;
    if (c0 < C) {
      const float4 g0 = *reinterpret_cast<const float4*>(w + c0), g1 = *reinterpret_cast<const float4*>(w + c0 + 4);
      const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
      float o[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = gg[i] * (v[rep][i] * rs);
      *reinterpret_cast<uint4*>(y + (size_t)row * C + c0) = pack8(o);
    }
  }
}

// act[row, 64 t + 32 b' ...] = silu(h1) * h2 over the fc_1|fc_2 GEMM's partial tiles; columns are [fc_1 x32 | fc_2 x32] blocks
__device__ __forceinline__ void swiglu_stage(const DecodePlan* p, const Phase& ph, int tid) {
  const int P = p->n_ctas, I = p->I;
  const int per_row = ph.tiles * 2 * 8;                 // (tile, 64-column block, 4-column group)
  const int total = p->n_seq * per_row;
  for (int e = blockIdx.x * DK_CTHREADS + tid; e < total; e += P * DK_CTHREADS) {
    const int row = e / per_row, rem = e % per_row;
    const int tile = rem >> 4, blk = (rem >> 3) & 1, j = (rem & 7) * 4;
    const int first = unit_owner(ph.units, tile * ph.nkb, P), last = unit_owner(ph.units, (tile + 1) * ph.nkb - 1, P);
    const float* base = p->ws + ((size_t)tile * ph.maxseg * 128 + row) * 128 + blk * 64 + j;
    float4 h1 = make_float4(0.f, 0.f, 0.f, 0.f), h2 = h1;
    for (int s = 0; s <= last - first; ++s) {
      const float4 a = __ldcg(reinterpret_cast<const float4*>(base + (size_t)s * 128 * 128));
      const float4 b = __ldcg(reinterpret_cast<const float4*>(base + (size_t)s * 128 * 128 + 32));
      h1.x += a.x; h1.y += a.y; h1.z += a.z; h1.w += a.w;
      h2.x += b.x; h2.y += b.y; h2.z += b.z; h2.w += b.w;
    }
    const float a1[4] = {h1.x, h1.y, h1.z, h1.w}, a2[4] = {h2.x, h2.y, h2.z, h2.w};
    float o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float u = bf16_round(a1[i]), g = bf16_round(a2[i]);
      o[i] = (u / (1.f + __expf(-u))) * g;
    }
    *reinterpret_cast<uint2*>(p->act + (size_t)row * I + tile * 64 + blk * 32 + j) = make_uint2(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]));
  }
}

// ---------------------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------------------
template <int NVB>
__global__ void __launch_bounds__(DK_THREADS, 1) decode_layers_kernel(const DecodePlan* __restrict__ plan) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* wring = smem;                                          // [W_STAGES][16 KB]
  uint8_t* aring = smem + DK_W_STAGES * DK_TILE_BYTES;            // [A_STAGES][16 KB]; the attention stage re-uses this region and beyond
  const DecodePlan* p = plan;
  const int att_bytes = dk_att_fixed_bytes(NVB) + p->qpk * (p->max_len + 20) * 4;
  const int region = att_bytes > DK_A_STAGES * DK_TILE_BYTES ? att_bytes : DK_A_STAGES * DK_TILE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(aring + ((region + 127) & ~127));
  uint64_t* w_full = bars;                        // [W_STAGES]
  uint64_t* w_empty = bars + DK_W_STAGES;         // [W_STAGES]
  uint64_t* a_full = bars + 2 * DK_W_STAGES;      // [A_STAGES]
  uint64_t* a_empty = a_full + DK_A_STAGES;       // [A_STAGES]
  uint64_t* tfull = a_empty + DK_A_STAGES;        // [2]
  uint64_t* tempty = tfull + 2;                   // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
  int* phase_go = reinterpret_cast<int*>(tmem_slot + 1);
  float* red = reinterpret_cast<float*>(tmem_slot + 4);           // [8]

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const int cta = blockIdx.x, P = p->n_ctas;
  const int n_layer = p->n_layer;
  const DecodeLayer* layers = reinterpret_cast<const DecodeLayer*>(p + 1);

  if (threadIdx.x == 0) {
    for (int i = 0; i < DK_W_STAGES; ++i) { mbar_init(&w_full[i], 1); mbar_init(&w_empty[i], 1); }
    for (int i = 0; i < DK_A_STAGES; ++i) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], DK_CWARPS); }
    *phase_go = 0;
    fence_barrier_init();
  }
  if (warp == DK_CWARPS + 2) { tmem_alloc(tmem_slot, 256); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

  if (warp == DK_CWARPS) {
    // ===================== weight producer: never waits for a grid barrier =====================
    if (lane == 0) {
      uint32_t it = 0;
      for (int l = 0; l < n_layer; ++l) {
        for (int g = 0; g < 4; ++g) {
          const Phase ph = get_phase(p, layers + l, g);
          const int u0 = unit_lo(ph.units, cta, P), u1 = unit_lo(ph.units, cta + 1, P);
          for (int u = u0; u < u1; ++u, ++it) {
            const int s = it % DK_W_STAGES;
            mbar_wait(&w_empty[s], ((it / DK_W_STAGES) & 1) ^ 1);
            mbar_expect_tx(&w_full[s], DK_TILE_BYTES);
            tma_load_2d(wring + s * DK_TILE_BYTES, ph.w, (u % ph.nkb) * 64, (u / ph.nkb) * 128, &w_full[s]);
          }
        }
      }
    }
  } else if (warp == DK_CWARPS + 1) {
    // ===================== activation producer: a stage's A boxes once its input exists =====================
    if (lane == 0) {
      uint32_t it = 0;
      for (int l = 0; l < n_layer; ++l) {
        for (int g = 0; g < 4; ++g) {
          const Phase ph = get_phase(p, layers + l, g);
          const int u0 = unit_lo(ph.units, cta, P), u1 = unit_lo(ph.units, cta + 1, P);
          if (u0 == u1) continue;
          const int need = l * 4 + g + 1;
          if (ld_volatile_shared(phase_go) < need) {
            const uint64_t t0 = globaltimer_ns();
            uint32_t spins = 0;
            while (ld_volatile_shared(phase_go) < need) {
              if (((++spins) & 0x3FF) == 0 && globaltimer_ns() - t0 > DH_HANG_GUARD_NS) dk_hang("activation producer", need, 0);
            }
          }
          fence_proxy_async_all();        // the activations were written through the generic proxy by other SMs
          for (int u = u0; u < u1; ++u, ++it) {
            const int s = it % DK_A_STAGES;
            mbar_wait(&a_empty[s], ((it / DK_A_STAGES) & 1) ^ 1);
            mbar_expect_tx(&a_full[s], DK_TILE_BYTES);
            tma_load_2d(aring + s * DK_TILE_BYTES, ph.a, (u % ph.nkb) * 64, 0, &a_full[s]);
          }
        }
      }
    }
  } else if (warp == DK_CWARPS + 2) {
    // ===================== MMA issuer (whole warp walks the loop, one elected lane issues) =====================
    constexpr uint32_t idesc = umma_idesc_bf16(128, 128, 0, 0);
    uint32_t it = 0, seg = 0;
    for (int l = 0; l < n_layer; ++l) {
      for (int g = 0; g < 4; ++g) {
        const Phase ph = get_phase(p, layers + l, g);
        const int u0 = unit_lo(ph.units, cta, P), u1 = unit_lo(ph.units, cta + 1, P);
        for (int u = u0; u < u1; ++u, ++it) {
          const int kb = u % ph.nkb;
          const bool first = (u == u0) || kb == 0, last = (u == u1 - 1) || kb == ph.nkb - 1;
          const int acc = seg & 1;
          if (first) {
            mbar_wait(&tempty[acc], ((seg >> 1) & 1) ^ 1);
            tc_fence_after();
          }
          const int ws = it % DK_W_STAGES, as = it % DK_A_STAGES;
          mbar_wait(&w_full[ws], (it / DK_W_STAGES) & 1);
          mbar_wait(&a_full[as], (it / DK_A_STAGES) & 1);
          tc_fence_after();
          const uint32_t sa = smem_u32(aring + as * DK_TILE_BYTES), sb = smem_u32(wring + ws * DK_TILE_BYTES);
          const uint32_t d_tmem = tmem_base + acc * 128;
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
              tc_mma_ss(d_tmem, umma_desc_sw128(sa + k * 32, 16, 1024), umma_desc_sw128(sb + k * 32, 16, 1024), idesc, (!first) || k != 0);
            tc_commit(&w_empty[ws]);
            tc_commit(&a_empty[as]);
            if (last) tc_commit(&tfull[acc]);
          }
          __syncwarp();
          if (last) ++seg;
        }
      }
    }
  } else {
    // ===================== compute warps =====================
    const int tid = threadIdx.x;
    unsigned long long* bar = p->bar;
    const unsigned long long gen = ld_acquire_u64(bar + 1);
    const unsigned long long nb_per_launch = (unsigned long long)n_layer * 8 - 1;     // layer 0 has no barrier in front of its QKV GEMM
    unsigned long long bcount = gen * nb_per_launch;                 // barriers completed before this launch
    uint32_t seg = 0;
    const int q4 = warp & 3, half = warp >> 2;

    // drain this CTA's partial tiles of one GEMM stage: TMEM -> fp32 slices in the workspace
    auto drain = [&](const Phase& ph, int stage_index) {
      if (tid == 0) {
        asm volatile("st.volatile.shared.s32 [%0], %1;" ::"r"(smem_u32(phase_go)), "r"(stage_index + 1) : "memory");
      }
      const int u0 = unit_lo(ph.units, cta, P), u1 = unit_lo(ph.units, cta + 1, P);
      int u = u0;
      while (u < u1) {
        const int tile = u / ph.nkb;
        const int end = (tile + 1) * ph.nkb < u1 ? (tile + 1) * ph.nkb : u1;
        const int acc = seg & 1;
        mbar_wait(&tfull[acc], (seg >> 1) & 1);
        tc_fence_after();
        uint32_t r0[32], r1[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(q4 * 32) << 16) + acc * 128 + half * 64;
        tmem_ld32(taddr, r0);
        tmem_ld32(taddr + 32, r1);
        tmem_wait_ld();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[acc]);
        const int row = q4 * 32 + lane;
        if (row < p->n_seq) {
          const int sidx = cta - unit_owner(ph.units, tile * ph.nkb, P);
          float4* dst = reinterpret_cast<float4*>(p->ws + (((size_t)tile * ph.maxseg + sidx) * 128 + row) * 128 + half * 64);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            dst[i] = make_float4(__uint_as_float(r0[4 * i]), __uint_as_float(r0[4 * i + 1]), __uint_as_float(r0[4 * i + 2]), __uint_as_float(r0[4 * i + 3]));
            dst[8 + i] = make_float4(__uint_as_float(r1[4 * i]), __uint_as_float(r1[4 * i + 1]), __uint_as_float(r1[4 * i + 2]), __uint_as_float(r1[4 * i + 3]));
          }
        }
        ++seg;
        u = end;
      }
    };

    for (int l = 0; l < n_layer; ++l) {
      const DecodeLayer* L = layers + l;
      const Phase g0 = get_phase(p, L, 0), g1 = get_phase(p, L, 1), g2 = get_phase(p, L, 2), g3 = get_phase(p, L, 3);
      // ---- QKV GEMM (input: n) ----
      if (l > 0) grid_barrier(bar, (++bcount) * P);      // layer 0's input was written before the launch
      drain(g0, l * 4 + 0);
      grid_barrier(bar, (++bcount) * P);
      // ---- attention over the KV cache: items (sequence, kv group) ----
      for (int item = cta; item < p->n_seq * p->G; item += P) {
        const int seq = item / p->G, g = item % p->G;
        const int pos = __ldcg(p->positions + seq);
        qkv_reduce_rope_cache(p, L, g0, seq, g, pos, tid);
        __threadfence_block();
        cbar_sync();
        attention_item<NVB>(p, L, aring, seq, g, pos + 1, tid);
      }
      grid_barrier(bar, (++bcount) * P);
      // ---- attention output projection ----
      drain(g1, l * 4 + 1);
      grid_barrier(bar, (++bcount) * P);
      if (cta < p->n_seq) resid_norm_row(p, g1, cta, p->x, p->x2, p->n, L->norm2, red, tid);
      grid_barrier(bar, (++bcount) * P);
      // ---- MLP ----
      drain(g2, l * 4 + 2);
      grid_barrier(bar, (++bcount) * P);
      swiglu_stage(p, g2, tid);
      grid_barrier(bar, (++bcount) * P);
      drain(g3, l * 4 + 3);
      grid_barrier(bar, (++bcount) * P);
      if (cta < p->n_seq) resid_norm_row(p, g3, cta, p->x2, p->x, l + 1 == n_layer ? p->hidden : p->n, L->norm_next, red, tid);
    }
    if (cta == 0 && tid == 0) {
      __threadfence();
      bar[1] = gen + 1;          // every CTA read `gen` before its first barrier arrival, and CTA 0 is past a barrier they all joined
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == DK_CWARPS + 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
}

static size_t decode_smem_bytes(int nvb, int qpk, int max_len) {
  const int att = dk_att_fixed_bytes(nvb) + qpk * (max_len + 20) * 4;
  const int region = att > DK_A_STAGES * DK_TILE_BYTES ? att : DK_A_STAGES * DK_TILE_BYTES;
  return (size_t)DK_W_STAGES * DK_TILE_BYTES + ((region + 127) & ~127) + 512 + 1024;
}

static int make_kmajor_map(CUtensorMap* map, const void* base, long long ld, long long rows, long long k) {
  uint64_t dims[2] = {(uint64_t)k, (uint64_t)rows};
  uint64_t strides[1] = {(uint64_t)ld * 2};
  uint32_t box[2] = {64, 128};
  return make_tensor_map(map, base, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

}  // namespace dh

using namespace dh;

extern "C" size_t dh_decode_plan_bytes(int n_layer) { return sizeof(DecodePlan) + (size_t)n_layer * sizeof(DecodeLayer); }

extern "C" size_t dh_decode_workspace_bytes(const dh_decode_desc* d) {
  if (!d) return 0;
  const int P = sm_count() > 0 ? sm_count() : 148;
  const int tiles[4] = {d->n_groups * (d->q_per_kv + 2) * 64 / 128, d->C / 128, 2 * d->I / 128, d->C / 128};
  const int nkb[4] = {d->C / 64, d->C / 64, d->C / 64, d->I / 64};
  size_t best = 0;
  for (int g = 0; g < 4; ++g) {
    const int units = tiles[g] * nkb[g];
    int maxseg = 1;
    for (int t = 0; t < tiles[g]; ++t) {
      const int c = unit_owner(units, (t + 1) * nkb[g] - 1, P) - unit_owner(units, t * nkb[g], P) + 1;
      if (c > maxseg) maxseg = c;
    }
    const size_t b = (size_t)tiles[g] * maxseg * 128 * 128 * sizeof(float);
    if (b > best) best = b;
  }
  return best;
}

extern "C" int dh_decode_plan_build(const dh_decode_desc* d, void* plan_host) {
  DH_CHECK(d && plan_host && d->layers, "dh_decode_plan_build: null argument");
  DH_CHECK(d->n_layer > 0 && d->n_layer <= DK_MAX_LAYERS, "dh_decode_plan_build: n_layer %d out of range", d->n_layer);
  DH_CHECK(d->n_seq > 0 && d->n_seq <= 128, "dh_decode_plan_build: n_seq %d must be in [1, 128]", d->n_seq);
  DH_CHECK(d->C % 128 == 0 && d->C <= 8192 && d->I % 64 == 0 && (2 * d->I) % 128 == 0, "dh_decode_plan_build: C %% 128, I %% 64 (C=%d I=%d)", d->C, d->I);
  DH_CHECK(d->q_per_kv >= 1 && d->q_per_kv <= DK_MAX_QPK && (d->n_groups * (d->q_per_kv + 2) * 64) % 128 == 0,
           "dh_decode_plan_build: q_per_kv %d / n_groups %d not supported", d->q_per_kv, d->n_groups);
  DH_CHECK(d->prefix_len >= 0 && d->prefix_len <= DK_MAX_PREFIX, "dh_decode_plan_build: prefix_len %d", d->prefix_len);
  DH_CHECK(d->x && d->x2 && d->n && d->qkv && d->y && d->act && d->hidden && d->ws && d->positions && d->rope_cos && d->rope_sin && d->barrier,
           "dh_decode_plan_build: null buffer");
  const int P = sm_count();
  DH_CHECK(P > 0, "dh_decode_plan_build: call dh_init first");
  int nvb = 4;
  if (decode_smem_bytes(4, d->q_per_kv, d->max_len) > 227 * 1024) nvb = 2;
  DH_CHECK(decode_smem_bytes(nvb, d->q_per_kv, d->max_len) <= 227 * 1024, "dh_decode_plan_build: max_len %d too long for the score buffer", d->max_len);
  DH_CHECK(d->ws_bytes >= dh_decode_workspace_bytes(d), "dh_decode_plan_build: workspace too small");
  DecodePlan* p = reinterpret_cast<DecodePlan*>(plan_host);
  memset(p, 0, dh_decode_plan_bytes(d->n_layer));
  const int qkv_out = d->n_groups * (d->q_per_kv + 2) * 64;
  if (make_kmajor_map(&p->a_n, d->n, d->C, d->n_seq, d->C)) return -3;
  if (make_kmajor_map(&p->a_y, d->y, d->C, d->n_seq, d->C)) return -3;
  if (make_kmajor_map(&p->a_act, d->act, d->I, d->n_seq, d->I)) return -3;
  p->n_layer = d->n_layer; p->n_seq = d->n_seq; p->C = d->C; p->I = d->I; p->G = d->n_groups; p->qpk = d->q_per_kv;
  p->max_len = d->max_len; p->prefix_len = d->prefix_len;
  p->n_ctas = P; p->nvb = nvb;
  p->eps = d->eps; p->scale = 1.0f / sqrtf((float)DK_HD);
  p->x = (float*)d->x; p->x2 = (float*)d->x2; p->n = (bf16*)d->n; p->qkv = (bf16*)d->qkv; p->y = (bf16*)d->y; p->act = (bf16*)d->act;
  p->hidden = (bf16*)d->hidden; p->ws = (float*)d->ws; p->positions = d->positions;
  p->rope_cos = (const bf16*)d->rope_cos; p->rope_sin = (const bf16*)d->rope_sin;
  p->bar = (unsigned long long*)d->barrier;
  const int tiles[4] = {qkv_out / 128, d->C / 128, 2 * d->I / 128, d->C / 128};
  const int nkb[4] = {d->C / 64, d->C / 64, d->C / 64, d->I / 64};
  for (int g = 0; g < 4; ++g) {
    const int units = tiles[g] * nkb[g];
    int maxseg = 1;
    for (int t = 0; t < tiles[g]; ++t) {
      const int c = unit_owner(units, (t + 1) * nkb[g] - 1, P) - unit_owner(units, t * nkb[g], P) + 1;
      if (c > maxseg) maxseg = c;
    }
    p->maxseg[g] = maxseg;
  }
  DecodeLayer* L = reinterpret_cast<DecodeLayer*>(p + 1);
  for (int l = 0; l < d->n_layer; ++l) {
    const dh_decode_layer* s = d->layers + l;
    DH_CHECK(s->wqkv && s->wproj && s->w12 && s->wmlp && s->norm2 && s->norm_next && s->cache_k && s->cache_v, "dh_decode_plan_build: layer %d: null pointer", l);
    if (make_kmajor_map(&L[l].wqkv, s->wqkv, d->C, qkv_out, d->C)) return -3;
    if (make_kmajor_map(&L[l].wproj, s->wproj, d->C, d->C, d->C)) return -3;
    if (make_kmajor_map(&L[l].w12, s->w12, d->C, 2 * d->I, d->C)) return -3;
    if (make_kmajor_map(&L[l].wmlp, s->wmlp, d->I, d->C, d->I)) return -3;
    L[l].norm2 = s->norm2; L[l].norm_next = s->norm_next;
    L[l].cache_k = (bf16*)s->cache_k; L[l].cache_v = (bf16*)s->cache_v;
    L[l].prefix_kv = (const bf16*)(d->prefix_len > 0 ? s->prefix_kv : nullptr);
    L[l].gate = L[l].prefix_kv ? s->gate : nullptr;
    DH_CHECK(!L[l].prefix_kv || L[l].gate, "dh_decode_plan_build: layer %d: prefix without gate", l);
  }
  return 0;
}

extern "C" int dh_decode_layers(const void* plan_dev, const void* plan_host, void* stream) {
  DH_CHECK(plan_dev && plan_host, "dh_decode_layers: null plan");
  const DecodePlan* h = reinterpret_cast<const DecodePlan*>(plan_host);
  const size_t smem = decode_smem_bytes(h->nvb, h->qpk, h->max_len);
  static bool attr = false;
  static int coop = -1;
  if (!attr) {
    DH_CUDA(cudaFuncSetAttribute(decode_layers_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    DH_CUDA(cudaFuncSetAttribute(decode_layers_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    const char* e = getenv("DH_DECODE_COOP");
    coop = (e && e[0] == '0') ? 0 : 1;
    attr = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(h->n_ctas);
  cfg.blockDim = dim3(DK_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeCooperative;      // all CTAs co-resident (the grid barriers need it)
  at[0].val.cooperative = 1;
  cfg.attrs = at;
  cfg.numAttrs = coop ? 1 : 0;
  const DecodePlan* pd = reinterpret_cast<const DecodePlan*>(plan_dev);
  if (h->nvb == 4) DH_CUDA(cudaLaunchKernelEx(&cfg, decode_layers_kernel<4>, pd));
  else DH_CUDA(cudaLaunchKernelEx(&cfg, decode_layers_kernel<2>, pd));
  return 0;
}
