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
//   * the KV cache is STREAMED like the weights: the activation producer's ring also carries the attention stage's K / V
//     chunks (64 keys: K 8 KB + V 8 KB per stage, 128-byte-swizzled TMA boxes of the cache rows).  Only keys written by
//     EARLIER launches are streamed (the new token's k / v stay in shared memory), so the loads have no dependency on this
//     step and start as soon as the QKV GEMM has freed the ring -- before the grid barrier in front of the attention.
//
// Warp roles (352 threads): warps 0-7 compute (barriers, partial-tile drain, consumers, attention), warp 8 weight TMA,
// warp 9 activation / KV-chunk TMA, warp 10 tcgen05.mma issue.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "decode_attn.cuh"
#include "../../include/dualhyp.h"

namespace dh {

constexpr int DK_CWARPS = 8;
constexpr int DK_CTHREADS = DK_CWARPS * 32;
constexpr int DK_THREADS = (DK_CWARPS + 3) * 32;
constexpr int DK_TILE_BYTES = 128 * 64 * 2;        // one [128 rows x 64 k] bf16 box = one ring stage
constexpr int DK_W_STAGES = 4;
constexpr int DK_A_STAGES = 7;
constexpr int DK_STG_BYTES = 32 * 1024;            // drain transpose staging (4 KB per compute warp) / attention scratch
constexpr int DK_KV_KEYS = KVC_KEYS;               // keys per streamed K / V chunk (K 8 KB + V 8 KB = one ring stage)
constexpr int DK_HD = 64;
constexpr int DK_MAX_QPK = 8;
constexpr int DK_MAX_PREFIX = 16;
constexpr int DK_MAX_LAYERS = 64;
constexpr int DK_TRACE_SLOTS = 24;

struct DecodeLayer {
  CUtensorMap wqkv, wproj, w12, wmlp;   // [N, K] bf16 K-major, box {64 k, 128 n}
  CUtensorMap kmap, vmap;               // KV cache as [n_seq * G * max_len, 64] bf16, box {64, DK_KV_KEYS keys}
  const float* norm2;                   // this layer's norm_2 weight
  const float* norm_next;               // the next layer's norm_1 weight (ln_f after the last layer)
  bf16* cache_k;
  bf16* cache_v;
  const bf16* prefix_kv;                // adapter prefix K/V rows [prefix_len, qkv_out] (NULL: no adapter in this layer)
  const float* gate;
  long long pad_[2];
};
static_assert(sizeof(DecodeLayer) % 64 == 0, "tensor maps inside the layer array must stay 64-byte aligned");
// the pointer members of a DecodeLayer: the kernel keeps a copy of these per layer in shared memory (with 200+ KB of shared
// memory per CTA the L1 is almost gone, so every read of the plan in global memory would be an L2 round trip)
struct LayerPtrs {
  const float* norm2;
  const float* norm_next;
  bf16* cache_k;
  bf16* cache_v;
  const bf16* prefix_kv;
  const float* gate;
};
static_assert(offsetof(DecodeLayer, norm2) + sizeof(LayerPtrs) <= sizeof(DecodeLayer), "LayerPtrs mirrors the tail of DecodeLayer");

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
  long long* trace;                     // optional [2 CTAs][n_layer][9] globaltimer stamps of the compute warps (NULL: off)
  long long pad_[2];
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
__device__ __forceinline__ void red_release_add_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ int ld_volatile_shared(const int* p) {
  int v;
  asm volatile("ld.volatile.shared.s32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
  return v;
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

[[noreturn]] static __device__ __noinline__ void dk_hang(const char* what, int a, int b) {
  printf("[dualhyp] decode step timed out in %s (block %d, %d / %d)\n", what, blockIdx.x, a, b);
  __trap();
  while (true) {}
}

// grid-wide barrier over the compute warps of all CTAs: every CTA arrives once on a monotonic counter.  bar.sync orders the
// CTA's writes before thread 0's release, and thread 0's acquire before the CTA's later reads (all cross-CTA reads in this
// kernel go to L2: ld.global.cg, TMA).
__device__ __forceinline__ void grid_barrier(unsigned long long* bar, unsigned long long target) {
  cbar_sync();
  if (threadIdx.x == 0) {
    red_release_add_u64(bar, 1ULL);
    if (ld_acquire_u64(bar) < target) {
      const uint64_t t0 = globaltimer_ns();
      uint32_t spins = 0;
      while (ld_acquire_u64(bar) < target) {
        if (((++spins) & 0xFF) == 0 && globaltimer_ns() - t0 > DH_HANG_GUARD_NS) dk_hang("grid barrier", (int)(target & 0x7fffffff), 0);
      }
    }
  }
  cbar_sync();
}

struct Phase {
  int tiles, nkb, maxseg, units;
  int P;            // CTAs that share this GEMM: min(grid, units), so that every one of them owns at least one unit
};
// `p` = the plan header (the kernel passes its shared-memory copy)
__device__ __forceinline__ Phase get_phase(const DecodePlan* p, int g) {
  Phase ph;
  if (g == 0) { ph.tiles = p->G * (p->qpk + 2) * DK_HD / 128; ph.nkb = p->C / 64; }
  else if (g == 1) { ph.tiles = p->C / 128; ph.nkb = p->C / 64; }
  else if (g == 2) { ph.tiles = 2 * p->I / 128; ph.nkb = p->C / 64; }
  else { ph.tiles = p->C / 128; ph.nkb = p->I / 64; }
  ph.maxseg = p->maxseg[g];
  ph.units = ph.tiles * ph.nkb;
  ph.P = ph.units < p->n_ctas ? ph.units : p->n_ctas;
  return ph;
}
// tensor maps live in GLOBAL memory (the plan as the host wrote it): weight map of GEMM g in layer L, activation map of GEMM g
__device__ __forceinline__ const CUtensorMap* weight_map(const DecodeLayer* L, int g) {
  return g == 0 ? &L->wqkv : g == 1 ? &L->wproj : g == 2 ? &L->w12 : &L->wmlp;
}
__device__ __forceinline__ const CUtensorMap* act_map(const DecodePlan* gp, int g) {
  return g == 1 ? &gp->a_y : g == 3 ? &gp->a_act : &gp->a_n;
}
// contiguous run of (tile, k-block) units of CTA c: [unit_lo(c), unit_lo(c + 1))
__device__ __host__ __forceinline__ int unit_lo(int units, int c, int P) { return c >= P ? units : (int)((long long)units * c / P); }
// the CTA whose run holds unit u
__device__ __host__ __forceinline__ int unit_owner(int units, int u, int P) { return (int)((((long long)u + 1) * P - 1) / units); }

// number of partial-tile slices of `tile` (the CTAs whose runs touch it are contiguous and each owns at least one unit)
__device__ __forceinline__ int tile_segments(const Phase& ph, int tile) {
  return unit_owner(ph.units, (tile + 1) * ph.nkb - 1, ph.P) - unit_owner(ph.units, tile * ph.nkb, ph.P) + 1;
}

__device__ __forceinline__ void f4_add(float4& a, const float4& b) { a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }

// Sum `nseg` partial-tile slices (16384 floats apart) at two float4 positions, slices in ascending order (deterministic); the
// loads of NB slices are issued together so that the reduction costs one L2 round trip, not nseg.
template <int NB = 4>
__device__ __forceinline__ void sum_slices2(const float* base, int off_b, int nseg, float4& lo, float4& hi) {
  lo = make_float4(0.f, 0.f, 0.f, 0.f);
  hi = lo;
  for (int s0 = 0; s0 < nseg; s0 += NB) {
    float4 a[NB], b[NB];
#pragma unroll
    for (int i = 0; i < NB; ++i) {
      a[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      b[i] = a[i];
      if (s0 + i < nseg) {
        a[i] = __ldcg(reinterpret_cast<const float4*>(base + (size_t)(s0 + i) * 128 * 128));
        b[i] = __ldcg(reinterpret_cast<const float4*>(base + (size_t)(s0 + i) * 128 * 128 + off_b));
      }
    }
#pragma unroll
    for (int i = 0; i < NB; ++i) { f4_add(lo, a[i]); f4_add(hi, b[i]); }       // absent slices add +0
  }
}

// ---------------------------------------------------------------------------------------------
// attention stage (one (sequence, kv group) item; 256 compute threads)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void dk_cp_async_16(uint32_t dst_smem, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst_smem), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void dk_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void dk_cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// attention scratch inside the 32 KB staging region
struct AttScratch {
  float part[DK_CWARPS][DK_MAX_QPK][DK_HD];     // per-warp un-normalised outputs over the warp's key chunks
  float pm[DK_CWARPS][DK_MAX_QPK];              // per-warp running max (log2 domain)
  float pl[DK_CWARPS][DK_MAX_QPK];              // per-warp sum of exp2
  float sout[DK_MAX_QPK][DK_HD];                // merged, normalised output of the cached keys + the new key
  float snew[DK_MAX_QPK];                       // the new key's score (log2 domain)
  float knew[DK_HD], vnew[DK_HD];               // the new token's k (after RoPE) and v, bf16-rounded
  bf16 sq[DK_MAX_QPK][DK_HD];                   // the new token's q heads of this group (after RoPE)
  bf16 spfx[DK_MAX_PREFIX][2 * DK_HD];          // adapter prefix rows [k 64 | v 64]
};
static_assert(sizeof(AttScratch) <= DK_STG_BYTES, "attention scratch must fit the staging region");

// Sum the QKV GEMM's partial tiles for row `seq`, columns of kv group g (qpk + 2 head slots), apply NeoX RoPE to the q and k
// slots at the row's position (the DH_EPI_QKV_ROPE math: bf16-rounded product, bf16 tables, pairs (j, j + 32)); the q heads
// stay in shared memory, the k / v slots go to the KV cache at (seq, pos) and to shared memory.  80 of the 256 threads work.
__device__ __forceinline__ void qkv_reduce_rope_cache(const DecodePlan* p, const LayerPtrs* L, const Phase& ph, AttScratch* sc, int seq, int g,
                                                      int pos, int tid) {
  const int slots = p->qpk + 2;
  for (int t = tid; t < slots * 8; t += DK_CTHREADS) {
    const int slot = t >> 3, j = (t & 7) * 4;
    const int col0 = (g * slots + slot) * DK_HD + j;
    const int tile = col0 >> 7, within = col0 & 127;
    const int first = unit_owner(ph.units, tile * ph.nkb, ph.P), last = unit_owner(ph.units, (tile + 1) * ph.nkb - 1, ph.P);
    float4 lo, hi;
    sum_slices2<4>(p->ws + ((size_t)tile * ph.maxseg * 128 + seq) * 128 + within, 32, last - first + 1, lo, hi);
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
    if (slot < p->qpk) {
      *reinterpret_cast<uint2*>(&sc->sq[slot][j]) = lo2;
      *reinterpret_cast<uint2*>(&sc->sq[slot][j + 32]) = hi2;
    } else {
      bf16* cd = (slot == p->qpk ? L->cache_k : L->cache_v) + (((long long)seq * p->G + g) * p->max_len + pos) * DK_HD + j;
      *reinterpret_cast<uint2*>(cd) = lo2;
      *reinterpret_cast<uint2*>(cd + 32) = hi2;
      float* dn = slot == p->qpk ? sc->knew : sc->vnew;
      const float2 a0 = unpack_bf16x2(lo2.x), a1 = unpack_bf16x2(lo2.y), b0 = unpack_bf16x2(hi2.x), b1 = unpack_bf16x2(hi2.y);
      dn[j] = a0.x; dn[j + 1] = a0.y; dn[j + 2] = a1.x; dn[j + 3] = a1.y;
      dn[j + 32] = b0.x; dn[j + 33] = b0.y; dn[j + 34] = b1.x; dn[j + 35] = b1.y;
    }
  }
}

// One (sequence, kv group) item: `pos` keys of earlier steps arrive through the activation ring as chunks of 64 keys (K tile
// at the stage base, V tile 8 KB above it; 128-byte rows, 16-byte chunks XOR-swizzled by row & 7); warp w owns chunks w,
// w + 8, ... and keeps a flash-style running (max, sum, output) for the group's q heads (rows of m16n8k16 tiles, native GQA);
// the warps' partial results and the new token's own key are merged in shared memory.  `ait` = ring index of the item's first
// chunk (advanced by the number of chunks).
__device__ __forceinline__ void attention_item(const DecodePlan* p, const LayerPtrs* L, uint8_t* aring, AttScratch* sc,
                                               uint64_t* a_full, uint64_t* a_empty, uint32_t* slot_done, uint32_t& ait, int seq, int g, int pos,
                                               int tid) {
  const int warp = tid >> 5, lane = tid & 31;
  const int qpk = p->qpk, G = p->G;
  const int prefix_len = L->prefix_kv ? p->prefix_len : 0;
  const float scale = p->scale;
  const float sl2 = scale * 1.4426950408889634f;
  const int r = lane >> 2, q = lane & 3;
  const long long ld_qkv = (long long)G * (qpk + 2) * DK_HD;

  qkv_reduce_rope_cache(p, L, get_phase(p, 0), sc, seq, g, pos, tid);
  if (tid < prefix_len * 16) {
    const int a = tid >> 4, c = tid & 15;
    dk_cp_async_16(smem_u32(&sc->spfx[a][c * 8]), L->prefix_kv + (long long)a * ld_qkv + (long long)(g * (qpk + 2) + qpk) * DK_HD + c * 8, 16);
  }
  dk_cp_async_commit();
  cbar_sync();                                  // sq / knew / vnew visible

  // q fragments (A operand rows = heads; rows >= qpk are zero)
  uint32_t qa[4][2];
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) {
    qa[ks][0] = r < qpk ? *reinterpret_cast<const uint32_t*>(&sc->sq[r][16 * ks + 2 * q]) : 0u;
    qa[ks][1] = r < qpk ? *reinterpret_cast<const uint32_t*>(&sc->sq[r][16 * ks + 8 + 2 * q]) : 0u;
  }
  KvcState st;
  st.init();
  const int nch = (pos + DK_KV_KEYS - 1) / DK_KV_KEYS;
  for (int c = warp; c < nch; c += DK_CWARPS) {
    const uint32_t it = ait + c;
    const int s = it % DK_A_STAGES;
    // A parity wait may look at most ONE phase ahead.  Chunk c shares its stage with chunk c - A_STAGES, which ANOTHER warp
    // consumes: wait for that warp's release through a monotonic per-stage counter (it cannot alias the way a parity does)
    // -- after it the phase in flight on a_full[s] is this chunk's.
    if (c >= DK_A_STAGES) {
      const uint32_t want = it - DK_A_STAGES + 1;
      if ((uint32_t)ld_volatile_shared(reinterpret_cast<const int*>(slot_done + s)) < want) {
        const uint64_t t0 = globaltimer_ns();
        uint32_t spins = 0;
        while ((uint32_t)ld_volatile_shared(reinterpret_cast<const int*>(slot_done + s)) < want) {
          if (((++spins) & 0x3FF) == 0 && globaltimer_ns() - t0 > DH_HANG_GUARD_NS) dk_hang("kv stage hand-over", (int)want, s);
        }
      }
    }
    mbar_wait(&a_full[s], (it / DK_A_STAGES) & 1);
    kvc_chunk_step(st, qa, smem_u32(aring + s * DK_TILE_BYTES), c * DK_KV_KEYS, pos, sl2, lane);
    __syncwarp();
    if (lane == 0) {
      asm volatile("st.volatile.shared.u32 [%0], %1;" ::"r"(smem_u32(slot_done + s)), "r"(it + 1) : "memory");
      mbar_arrive(&a_empty[s]);                        // the stage may be refilled
    }
  }
  ait += nch;
  st.l += __shfl_xor_sync(0xffffffffu, st.l, 1);
  st.l += __shfl_xor_sync(0xffffffffu, st.l, 2);
#pragma unroll
  for (int dt = 0; dt < 8; ++dt) *reinterpret_cast<float2*>(&sc->part[warp][r][8 * dt + 2 * q]) = make_float2(st.acc[dt][0], st.acc[dt][1]);
  if (q == 0) { sc->pm[warp][r] = st.m; sc->pl[warp][r] = st.l; }
  // the new token's own key: score per head (warp = head)
  float q0 = 0.f, q1 = 0.f, gate_h = 0.f;
  if (warp < qpk) {
    q0 = __bfloat162float(sc->sq[warp][lane]);
    q1 = __bfloat162float(sc->sq[warp][lane + 32]);
    const float d = warp_sum(q0 * sc->knew[lane] + q1 * sc->knew[lane + 32]);
    if (lane == 0) sc->snew[warp] = d * sl2;
    q0 *= scale;
    q1 *= scale;
    if (prefix_len > 0) gate_h = L->gate[g * qpk + warp];
  }
  dk_cp_async_wait<0>();
  cbar_sync();
  // ---- merge the warps' partial results and the new key ----
  for (int i = tid; i < DK_MAX_QPK * DK_HD; i += DK_CTHREADS) {
    const int h = i >> 6, d = i & 63;
    if (h < qpk) {
      const float sn = sc->snew[h];
      float M = sn;
#pragma unroll
      for (int w = 0; w < DK_CWARPS; ++w) M = fmaxf(M, sc->pm[w][h]);
      const float en = exp2f(sn - M);
      float num = en * sc->vnew[d], den = en;
#pragma unroll
      for (int w = 0; w < DK_CWARPS; ++w) {
        const float e = exp2f(sc->pm[w][h] - M);
        num += e * sc->part[w][h][d];
        den += e * sc->pl[w][h];
      }
      sc->sout[h][d] = num / den;
    }
  }
  cbar_sync();
  if (warp < qpk) {
    float pfx0 = 0.f, pfx1 = 0.f;
    if (prefix_len > 0) {
      float ps[DK_MAX_PREFIX];
      float pmx = -INFINITY;
#pragma unroll
      for (int a = 0; a < DK_MAX_PREFIX; ++a) {
        ps[a] = -INFINITY;
        if (a < prefix_len) {
          ps[a] = warp_sum(q0 * __bfloat162float(sc->spfx[a][lane]) + q1 * __bfloat162float(sc->spfx[a][lane + 32]));
          pmx = fmaxf(pmx, ps[a]);
        }
      }
      float psum = 0.f;
#pragma unroll
      for (int a = 0; a < DK_MAX_PREFIX; ++a) {
        ps[a] = (a < prefix_len) ? __expf(ps[a] - pmx) : 0.f;
        psum += ps[a];
      }
      const float gs = gate_h / psum;
#pragma unroll
      for (int a = 0; a < DK_MAX_PREFIX; ++a) {
        if (a < prefix_len) {
          pfx0 += gs * ps[a] * __bfloat162float(sc->spfx[a][64 + lane]);
          pfx1 += gs * ps[a] * __bfloat162float(sc->spfx[a][96 + lane]);
        }
      }
    }
    const int h = g * qpk + warp;
    const float o0 = sc->sout[warp][lane] + pfx0, o1 = sc->sout[warp][lane + 32] + pfx1;
    bf16* orow = p->y + (long long)seq * (G * qpk) * DK_HD + h * DK_HD;
    orow[lane] = __float2bfloat16_rn(o0);
    orow[lane + 32] = __float2bfloat16_rn(o1);
  }
  cbar_sync();      // the scratch is re-used by the next item
}

// ---------------------------------------------------------------------------------------------
// consumer stages
// ---------------------------------------------------------------------------------------------
// row `row`: x_out = resid + bf16(sum of the GEMM's partial tiles), y = rmsnorm(x_out) * w  (256 threads, C <= 4096)
__device__ __forceinline__ void resid_norm_row(const DecodePlan* p, int gemm, int row, const float* resid, float* x_out, bf16* y,
                                               const float* w, float* red, int tid) {
  const Phase ph = get_phase(p, gemm);
  const int C = p->C;
  float v[2][8];
  float ss = 0.f;
#pragma unroll
  for (int rep = 0; rep < 2; ++rep) {
    const int c0 = rep * 2048 + tid * 8;
    if (c0 < C) {
      const int tile = c0 >> 7, within = c0 & 127;
      float4 s0, s1;
      sum_slices2(p->ws + ((size_t)tile * ph.maxseg * 128 + row) * 128 + within, 4, tile_segments(ph, tile), s0, s1);
      const float acc[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
      const float4 r0 = __ldcg(reinterpret_cast<const float4*>(resid + (size_t)row * C + c0));
      const float4 r1 = __ldcg(reinterpret_cast<const float4*>(resid + (size_t)row * C + c0 + 4));
      const float rr[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        v[rep][i] = rr[i] + bf16_round(acc[i]);
        ss += v[rep][i] * v[rep][i];
      }
      __stcg(reinterpret_cast<float4*>(x_out + (size_t)row * C + c0), make_float4(v[rep][0], v[rep][1], v[rep][2], v[rep][3]));
      __stcg(reinterpret_cast<float4*>(x_out + (size_t)row * C + c0 + 4), make_float4(v[rep][4], v[rep][5], v[rep][6], v[rep][7]));
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
  for (int rep = 0; rep < 2; ++rep) {
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
__device__ __forceinline__ void swiglu_stage(const DecodePlan* p, int tid) {
  const Phase ph = get_phase(p, 2);
  const int P = p->n_ctas, I = p->I;
  const int per_row = ph.tiles * 2 * 8;                 // (tile, 64-column block, 4-column group)
  const int total = p->n_seq * per_row;
  const int stride = P * DK_CTHREADS;
  for (int e0 = blockIdx.x * DK_CTHREADS + tid; e0 < total; e0 += 2 * stride) {
    float4 h1[2], h2[2];
    int rows[2], cols[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {                       // two elements in flight per thread
      const int e = e0 + k * stride;
      rows[k] = -1;
      if (e < total) {
        const int row = e / per_row, rem = e % per_row;
        const int tile = rem >> 4, blk = (rem >> 3) & 1, j = (rem & 7) * 4;
        rows[k] = row;
        cols[k] = tile * 64 + blk * 32 + j;
        sum_slices2(p->ws + ((size_t)tile * ph.maxseg * 128 + row) * 128 + blk * 64 + j, 32, tile_segments(ph, tile), h1[k], h2[k]);
      }
    }
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      if (rows[k] >= 0) {
        const float a1[4] = {h1[k].x, h1[k].y, h1[k].z, h1[k].w}, a2[4] = {h2[k].x, h2[k].y, h2[k].z, h2[k].w};
        float o[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float u = bf16_round(a1[i]), g = bf16_round(a2[i]);
          float th;
      asm("tanh.approx.f32 %0, %1;" : "=f"(th) : "f"(0.5f * u));        // sigmoid(u) = 0.5 + 0.5 tanh(u / 2), as the GEMM epilogues
      o[i] = u * fmaf(0.5f, th, 0.5f) * g;
        }
        *reinterpret_cast<uint2*>(p->act + (size_t)rows[k] * I + cols[k]) = make_uint2(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]));
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------------------
template <bool TRACE>
__global__ void __launch_bounds__(DK_THREADS, 1) decode_layers_kernel(const DecodePlan* __restrict__ plan) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* wring = smem;                                          // [W_STAGES][16 KB] weight tiles
  uint8_t* aring = smem + DK_W_STAGES * DK_TILE_BYTES;            // [A_STAGES][16 KB] activation boxes / K+V chunks
  uint8_t* stg = aring + DK_A_STAGES * DK_TILE_BYTES;             // 32 KB: drain transpose staging / attention scratch
  const DecodePlan* gp = plan;                                    // global copy: the tensor maps TMA reads
  uint64_t* bars = reinterpret_cast<uint64_t*>(stg + DK_STG_BYTES);
  uint64_t* w_full = bars;                        // [W_STAGES]
  uint64_t* w_empty = bars + DK_W_STAGES;         // [W_STAGES]
  uint64_t* a_full = bars + 2 * DK_W_STAGES;      // [A_STAGES]
  uint64_t* a_empty = a_full + DK_A_STAGES;       // [A_STAGES]
  uint64_t* tfull = a_empty + DK_A_STAGES;        // [2]
  uint64_t* tempty = tfull + 2;                   // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
  int* phase_go = reinterpret_cast<int*>(tmem_slot + 1);
  float* red = reinterpret_cast<float*>(tmem_slot + 4);           // [8]
  uint32_t* slot_done = reinterpret_cast<uint32_t*>(red + 8);     // [A_STAGES] ring index + 1 of the last K/V chunk released per stage
  // shared-memory copies of the plan header and of every layer's pointers (see LayerPtrs)
  DecodePlan* sp = reinterpret_cast<DecodePlan*>(reinterpret_cast<uint8_t*>(bars) + 512);
  LayerPtrs* slp = reinterpret_cast<LayerPtrs*>(sp + 1);
  const DecodeLayer* layers = reinterpret_cast<const DecodeLayer*>(gp + 1);
  {
    const int n_layer_g = gp->n_layer;
    for (int i = threadIdx.x; i < (int)(sizeof(DecodePlan) / 16); i += DK_THREADS)
      reinterpret_cast<uint4*>(sp)[i] = reinterpret_cast<const uint4*>(gp)[i];
    for (int i = threadIdx.x; i < n_layer_g * 6; i += DK_THREADS)
      reinterpret_cast<uint64_t*>(slp)[i] = *reinterpret_cast<const uint64_t*>(reinterpret_cast<const uint8_t*>(layers + i / 6) + offsetof(DecodeLayer, norm2) + (i % 6) * 8);
  }
  const DecodePlan* p = sp;

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const int cta = blockIdx.x;

  if (threadIdx.x == 0) {
    for (int i = 0; i < DK_W_STAGES; ++i) { mbar_init(&w_full[i], 1); mbar_init(&w_empty[i], 1); }
    for (int i = 0; i < DK_A_STAGES; ++i) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], DK_CWARPS); }
    *phase_go = 0;
    for (int i = 0; i < DK_A_STAGES; ++i) slot_done[i] = 0;
    fence_barrier_init();
  }
  if (warp == DK_CWARPS + 2) { tmem_alloc(tmem_slot, 256); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  const int P = p->n_ctas;
  const int n_layer = p->n_layer;
  const int n_items = p->n_seq * p->G;

  // K/V chunks this CTA streams per layer (its attention items are the same in every layer)
  int my_chunks = 0;
  for (int item = cta; item < n_items; item += P) my_chunks += (__ldg(p->positions + item / p->G) + DK_KV_KEYS - 1) / DK_KV_KEYS;

  if (warp == DK_CWARPS) {
    // ===================== weight producer: never waits for a grid barrier =====================
    if (lane == 0) {
      uint32_t it = 0;
      for (int l = 0; l < n_layer; ++l) {
        for (int g = 0; g < 4; ++g) {
          const Phase ph = get_phase(p, g);
          const CUtensorMap* wmap = weight_map(layers + l, g);
          const int u0 = unit_lo(ph.units, cta, ph.P), u1 = unit_lo(ph.units, cta + 1, ph.P);
          for (int u = u0; u < u1; ++u, ++it) {
            const int s = it % DK_W_STAGES;
            mbar_wait(&w_empty[s], ((it / DK_W_STAGES) & 1) ^ 1);
            mbar_expect_tx(&w_full[s], DK_TILE_BYTES);
            tma_load_2d(wring + s * DK_TILE_BYTES, wmap, (u % ph.nkb) * 64, (u / ph.nkb) * 128, &w_full[s]);
          }
        }
      }
    }
  } else if (warp == DK_CWARPS + 1) {
    // ===================== activation producer: a stage's A boxes once its input exists; K/V chunks as soon as the ring is free ==========
    if (lane == 0) {
      uint32_t it = 0;
      for (int l = 0; l < n_layer; ++l) {
        const DecodeLayer* L = layers + l;
        for (int g = 0; g < 4; ++g) {
          const Phase ph = get_phase(p, g);
          const CUtensorMap* amap = act_map(gp, g);
          const int u0 = unit_lo(ph.units, cta, ph.P), u1 = unit_lo(ph.units, cta + 1, ph.P);
          if (u0 < u1) {
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
              tma_load_2d(aring + s * DK_TILE_BYTES, amap, (u % ph.nkb) * 64, 0, &a_full[s]);
            }
          }
          if (g == 0) {
            // the cached keys of this CTA's attention items (written by earlier launches: no dependency on this step)
            for (int item = cta; item < n_items; item += P) {
              const int seq = item / p->G, kg = item % p->G;
              const int pos = __ldg(p->positions + seq);
              const int row0 = (seq * p->G + kg) * p->max_len;
              for (int c = 0; c * DK_KV_KEYS < pos; ++c, ++it) {
                const int s = it % DK_A_STAGES;
                mbar_wait(&a_empty[s], ((it / DK_A_STAGES) & 1) ^ 1);
                mbar_expect_tx(&a_full[s], DK_TILE_BYTES);
                tma_load_2d(aring + s * DK_TILE_BYTES, &L->kmap, 0, row0 + c * DK_KV_KEYS, &a_full[s]);
                tma_load_2d(aring + s * DK_TILE_BYTES + DK_KV_KEYS * 128, &L->vmap, 0, row0 + c * DK_KV_KEYS, &a_full[s]);
              }
            }
          }
        }
      }
    }
  } else if (warp == DK_CWARPS + 2) {
    // ===================== MMA issuer (whole warp walks the loop, one elected lane issues) =====================
    constexpr uint32_t idesc = umma_idesc_bf16(128, 128, 0, 0);
    uint32_t wit = 0, ait = 0, seg = 0;
    for (int l = 0; l < n_layer; ++l) {
      for (int g = 0; g < 4; ++g) {
        const Phase ph = get_phase(p, g);
        const int u0 = unit_lo(ph.units, cta, ph.P), u1 = unit_lo(ph.units, cta + 1, ph.P);
        if (u0 < u1) {
          // A parity wait may look at most ONE phase ahead, and the attention stage's K/V chunks cycle through the same ring
          // slots between the QKV and the proj GEMM: do not touch a_full before the compute warps have opened this stage
          // (by then every earlier use of the ring has completed).
          const int need = l * 4 + g + 1;
          if (ld_volatile_shared(phase_go) < need) {
            const uint64_t t0 = globaltimer_ns();
            uint32_t spins = 0;
            while (ld_volatile_shared(phase_go) < need) {
              if (((++spins) & 0x3FF) == 0 && globaltimer_ns() - t0 > DH_HANG_GUARD_NS) dk_hang("mma issuer", need, 0);
            }
          }
        }
        for (int u = u0; u < u1; ++u, ++wit, ++ait) {
          const int kb = u % ph.nkb;
          const bool first = (u == u0) || kb == 0, last = (u == u1 - 1) || kb == ph.nkb - 1;
          const int acc = seg & 1;
          if (first) {
            mbar_wait(&tempty[acc], ((seg >> 1) & 1) ^ 1);
            tc_fence_after();
          }
          const int ws = wit % DK_W_STAGES, as = ait % DK_A_STAGES;
          mbar_wait(&w_full[ws], (wit / DK_W_STAGES) & 1);
          mbar_wait(&a_full[as], (ait / DK_A_STAGES) & 1);
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
        if (g == 0) ait += my_chunks;          // the attention stage's K/V chunks pass through the same ring
      }
    }
  } else {
    // ===================== compute warps =====================
    const int tid = threadIdx.x;
    unsigned long long* bar = p->bar;
    const unsigned long long gen = ld_acquire_u64(bar + 1);
    const unsigned long long nb_per_launch = (unsigned long long)n_layer * 8 - 1;     // layer 0 has no barrier in front of its QKV GEMM
    unsigned long long bcount = gen * nb_per_launch;                 // barriers completed before this launch
    uint32_t seg = 0, ait = 0;
    const int q4 = warp & 3, half = warp >> 2;
    long long t_first_full = 0;
    AttScratch* sc = reinterpret_cast<AttScratch*>(stg);
    float* stgw = reinterpret_cast<float*>(stg) + warp * 1024;       // this warp's [32][32] transpose tile

    // drain this CTA's partial tiles of one GEMM stage: TMEM -> (shared-memory transpose) -> fp32 slices in the workspace, so
    // that every store instruction of a warp writes 128 contiguous bytes of one row
    auto drain = [&](const Phase& ph, int stage_index) {
      if (tid == 0) {
        asm volatile("st.volatile.shared.s32 [%0], %1;" ::"r"(smem_u32(phase_go)), "r"(stage_index + 1) : "memory");
      }
      const int u0 = unit_lo(ph.units, cta, ph.P), u1 = unit_lo(ph.units, cta + 1, ph.P);
      ait += u1 - u0;
      int u = u0;
      while (u < u1) {
        const int tile = u / ph.nkb;
        const int end = (tile + 1) * ph.nkb < u1 ? (tile + 1) * ph.nkb : u1;
        const int acc = seg & 1;
        mbar_wait(&tfull[acc], (seg >> 1) & 1);
        tc_fence_after();
        if (TRACE && u == u0 && tid == 0) t_first_full = (long long)globaltimer_ns();
        const int sidx = cta - unit_owner(ph.units, tile * ph.nkb, ph.P);
        float* dst = p->ws + (((size_t)tile * ph.maxseg + sidx) * 128 + q4 * 32) * 128 + half * 64 + lane;
        const int nrow = p->n_seq - q4 * 32;                       // valid rows of this warp's 32-row block
#pragma unroll
        for (int h2 = 0; h2 < 2; ++h2) {
          uint32_t rg[32];
          tmem_ld32(tmem_base + ((uint32_t)(q4 * 32) << 16) + acc * 128 + half * 64 + h2 * 32, rg);
          tmem_wait_ld();
          if (h2 == 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[acc]);              // the accumulator may be overwritten
          }
#pragma unroll
          for (int i = 0; i < 32; ++i) stgw[lane * 32 + (i ^ lane)] = __uint_as_float(rg[i]);
          __syncwarp();
#pragma unroll 8
          for (int rr = 0; rr < 32; ++rr) {
            if (rr < nrow) __stcg(dst + (size_t)rr * 128 + h2 * 32, stgw[rr * 32 + (lane ^ rr)]);
          }
          __syncwarp();
        }
        ++seg;
        u = end;
      }
    };

    // developer trace (DH_DECODE_TRACE=1): [3 CTAs][n_layer][24] stage stamps, then [8 barriers][P][2] arrive / release of layer 5
    long long* trace = TRACE ? p->trace : nullptr;
    long long* tr_bar = trace ? trace + (size_t)3 * n_layer * DK_TRACE_SLOTS : nullptr;
    if (trace && !(cta == 0 || cta == P / 2 || cta == P - 1)) trace = nullptr;
    if (trace) trace += (size_t)(cta == 0 ? 0 : cta == P - 1 ? 2 : 1) * n_layer * DK_TRACE_SLOTS;
    auto stamp = [&](int l, int i) {
      if (TRACE && trace && tid == 0) trace[l * DK_TRACE_SLOTS + i] = (long long)globaltimer_ns();
    };
    auto gbar = [&](int l, int b) {
      const bool rec = TRACE && tr_bar && tid == 0 && l == (n_layer > 5 ? 5 : n_layer - 1);
      if (rec) tr_bar[((size_t)b * P + cta) * 2] = (long long)globaltimer_ns();
      grid_barrier(bar, (++bcount) * P);
      if (rec) tr_bar[((size_t)b * P + cta) * 2 + 1] = (long long)globaltimer_ns();
    };
    for (int l = 0; l < n_layer; ++l) {
      const LayerPtrs* L = slp + l;
      stamp(l, 0);
      // ---- QKV GEMM (input: n) ----
      if (l > 0) gbar(l, 0);      // layer 0's input was written before the launch
      stamp(l, 1);
      drain(get_phase(p, 0), l * 4 + 0);
      if (TRACE && trace && tid == 0) trace[l * DK_TRACE_SLOTS + 2] = t_first_full;
      stamp(l, 3);
      gbar(l, 1);
      stamp(l, 4);
      // ---- attention over the KV cache: items (sequence, kv group) ----
      for (int item = cta; item < n_items; item += P) {
        const int seq = item / p->G, g = item % p->G;
        attention_item(p, L, aring, sc, a_full, a_empty, slot_done, ait, seq, g, __ldg(p->positions + seq), tid);
        if (item == cta) stamp(l, 5);
      }
      stamp(l, 6);
      gbar(l, 2);
      stamp(l, 7);
      // ---- attention output projection ----
      drain(get_phase(p, 1), l * 4 + 1);
      if (TRACE && trace && tid == 0) trace[l * DK_TRACE_SLOTS + 8] = t_first_full;
      stamp(l, 9);
      gbar(l, 3);
      stamp(l, 10);
      if (cta < p->n_seq) resid_norm_row(p, 1, cta, p->x, p->x2, p->n, L->norm2, red, tid);
      stamp(l, 11);
      gbar(l, 4);
      stamp(l, 12);
      // ---- MLP ----
      drain(get_phase(p, 2), l * 4 + 2);
      if (TRACE && trace && tid == 0) trace[l * DK_TRACE_SLOTS + 13] = t_first_full;
      stamp(l, 14);
      gbar(l, 5);
      stamp(l, 15);
      swiglu_stage(p, tid);
      stamp(l, 16);
      gbar(l, 6);
      stamp(l, 17);
      drain(get_phase(p, 3), l * 4 + 3);
      if (TRACE && trace && tid == 0) trace[l * DK_TRACE_SLOTS + 18] = t_first_full;
      stamp(l, 19);
      gbar(l, 7);
      stamp(l, 20);
      if (cta < p->n_seq) resid_norm_row(p, 3, cta, p->x2, p->x, l + 1 == n_layer ? p->hidden : p->n, L->norm_next, red, tid);
      stamp(l, 21);
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

static size_t decode_smem_bytes() {
  return (size_t)(DK_W_STAGES + DK_A_STAGES) * DK_TILE_BYTES + DK_STG_BYTES + 512 + sizeof(DecodePlan) + DK_MAX_LAYERS * sizeof(LayerPtrs) + 1024;
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
    const int Pg = units < P ? units : P;
    int maxseg = 1;
    for (int t = 0; t < tiles[g]; ++t) {
      const int c = unit_owner(units, (t + 1) * nkb[g] - 1, Pg) - unit_owner(units, t * nkb[g], Pg) + 1;
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
  DH_CHECK(d->C % 128 == 0 && d->C <= 4096 && d->I % 64 == 0 && (2 * d->I) % 128 == 0, "dh_decode_plan_build: C %% 128, I %% 64 (C=%d I=%d)", d->C, d->I);
  DH_CHECK(d->q_per_kv >= 1 && d->q_per_kv <= DK_MAX_QPK && (d->n_groups * (d->q_per_kv + 2) * 64) % 128 == 0,
           "dh_decode_plan_build: q_per_kv %d / n_groups %d not supported", d->q_per_kv, d->n_groups);
  DH_CHECK(d->prefix_len >= 0 && d->prefix_len <= DK_MAX_PREFIX, "dh_decode_plan_build: prefix_len %d", d->prefix_len);
  DH_CHECK(d->x && d->x2 && d->n && d->qkv && d->y && d->act && d->hidden && d->ws && d->positions && d->rope_cos && d->rope_sin && d->barrier,
           "dh_decode_plan_build: null buffer");
  const int P = sm_count();
  DH_CHECK(P > 0, "dh_decode_plan_build: call dh_init first");
  DH_CHECK(d->max_len > 0 && d->max_len % DK_KV_KEYS == 0, "dh_decode_plan_build: max_len %d must be a multiple of %d (K/V chunks)", d->max_len, DK_KV_KEYS);
  static_assert((DK_W_STAGES + DK_A_STAGES) * DK_TILE_BYTES + DK_STG_BYTES + 512 + sizeof(DecodePlan) + DK_MAX_LAYERS * sizeof(LayerPtrs) + 1024 <= 227 * 1024,
                "shared memory budget");
  DH_CHECK(d->ws_bytes >= dh_decode_workspace_bytes(d), "dh_decode_plan_build: workspace too small");
  DecodePlan* p = reinterpret_cast<DecodePlan*>(plan_host);
  memset(p, 0, dh_decode_plan_bytes(d->n_layer));
  const int qkv_out = d->n_groups * (d->q_per_kv + 2) * 64;
  if (make_kmajor_map(&p->a_n, d->n, d->C, d->n_seq, d->C)) return -3;
  if (make_kmajor_map(&p->a_y, d->y, d->C, d->n_seq, d->C)) return -3;
  if (make_kmajor_map(&p->a_act, d->act, d->I, d->n_seq, d->I)) return -3;
  p->n_layer = d->n_layer; p->n_seq = d->n_seq; p->C = d->C; p->I = d->I; p->G = d->n_groups; p->qpk = d->q_per_kv;
  p->max_len = d->max_len; p->prefix_len = d->prefix_len;
  p->n_ctas = P; p->nvb = 0;
  p->eps = d->eps; p->scale = 1.0f / sqrtf((float)DK_HD);
  p->x = (float*)d->x; p->x2 = (float*)d->x2; p->n = (bf16*)d->n; p->qkv = (bf16*)d->qkv; p->y = (bf16*)d->y; p->act = (bf16*)d->act;
  p->hidden = (bf16*)d->hidden; p->ws = (float*)d->ws; p->positions = d->positions;
  p->rope_cos = (const bf16*)d->rope_cos; p->rope_sin = (const bf16*)d->rope_sin;
  p->bar = (unsigned long long*)d->barrier;
  p->trace = (long long*)d->trace;
  const int tiles[4] = {qkv_out / 128, d->C / 128, 2 * d->I / 128, d->C / 128};
  const int nkb[4] = {d->C / 64, d->C / 64, d->C / 64, d->I / 64};
  for (int g = 0; g < 4; ++g) {
    const int units = tiles[g] * nkb[g];
    const int Pg = units < P ? units : P;
    int maxseg = 1;
    for (int t = 0; t < tiles[g]; ++t) {
      const int c = unit_owner(units, (t + 1) * nkb[g] - 1, Pg) - unit_owner(units, t * nkb[g], Pg) + 1;
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
    if (make_kv_cache_map(&L[l].kmap, s->cache_k, (long long)d->n_seq * d->n_groups * d->max_len)) return -3;
    if (make_kv_cache_map(&L[l].vmap, s->cache_v, (long long)d->n_seq * d->n_groups * d->max_len)) return -3;
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
  const size_t smem = decode_smem_bytes();
  static bool attr = false;
  static int coop = -1;
  if (!attr) {
    DH_CUDA(cudaFuncSetAttribute(decode_layers_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    DH_CUDA(cudaFuncSetAttribute(decode_layers_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
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
  if (h->trace) DH_CUDA(cudaLaunchKernelEx(&cfg, decode_layers_kernel<true>, pd));
  else DH_CUDA(cudaLaunchKernelEx(&cfg, decode_layers_kernel<false>, pd));
  return 0;
}
