// Stand-alone streamed decode attention (the default behind dh_attention_decode): item = (sequence, kv group).  ONE resident
// block per SM (15 compute warps + a producer warp = 4 warps per SM sub-partition, so each may hold 128 registers; a 17th warp
// would cap all of them at 96), each block keeping one kv group -- its adapter prefix rows are staged once -- and walking its share
// of the sequences.  The producer warp streams the cached K / V rows of the block's items as ONE stream of 32-key chunks through a
// 15-stage TMA ring (chunk n -> stage n % 15 -> compute warp n % 15): it runs ahead into the next item while the compute warps merge
// the current one, and the chunks that hold only keys of earlier token steps are requested BEFORE griddepcontrol.wait, so they
// overlap the tail of the QKV kernel.  Each compute warp keeps a flash-style running state for the group's q heads on mma.sync tiles
// (decode_attn.cuh); the adapter prefix (own softmax and gate, ger/adapter.py:116-146) is one more chunk, taken by the warp that
// has the fewest chunks of the item; the states are merged in shared memory and stored straight to the output row.
// Measured (scripts/attn_decode_probe.py, 22 calls per graph on cold caches, 256 sequences of 394..545 keys): 42.1 us per call as
// one block of 8 warps per item -> 28.8 us (4.3 TB/s); what each step bought is in DESIGN.md 3.3.
// DH_DEC_ATTN_DBG selects ablation builds of the same kernel: 1 no chunk math, 2 no loads, 3 cycle / globaltimer trace
// (dh_debug_ds_trace), 4 no merge, 5 no prefix chunk.
#include <math.h>
#include <stdlib.h>

#include "decode_attn.cuh"
#include "../../include/dualhyp.h"

namespace dh {

constexpr int DS_CWARPS = 15;           // == DS_STAGES: every ring stage is consumed by ONE warp (see the chunk loop)
constexpr int DS_CTHREADS = DS_CWARPS * 32;
constexpr int DS_THREADS = DS_CTHREADS + 32;
constexpr int DS_STAGES = 15;
constexpr int DS_KEYS = 32;             // keys per chunk: K 4 KB + V 4 KB per stage.  A warp's chunk is one long dependent chain (ldmatrix ->
                                        // mma -> softmax -> mma, ~5 us per 64 keys), so latency is hidden by warps, not by deeper chunks: the first
                                        // build (3 warps x 64 keys, 3 blocks per SM = 9 compute warps per SM) measured 48 us per layer at 256 sequences
constexpr int DS_STAGE_BYTES = 2 * DS_KEYS * 128;
constexpr int DS_MAX_QPK = 8;
constexpr int DS_MAX_PREFIX = DS_KEYS;     // the prefix is one chunk of its own

struct DsScratch {
  uint8_t pfx[DS_STAGE_BYTES];            // the kv group's adapter prefix rows laid out like a ring stage (K tile | V tile, swizzled rows)
  float part[DS_CWARPS][DS_MAX_QPK][64];
  float pm[DS_CWARPS][DS_MAX_QPK];
  float pl[DS_CWARPS][DS_MAX_QPK];
  float spo[DS_MAX_QPK][64];              // gate * softmax(prefix scores) @ prefix V of the current item
  uint64_t full[DS_STAGES], empty[DS_STAGES];
};
constexpr int DS_SMEM = DS_STAGES * DS_STAGE_BYTES + (int)sizeof(DsScratch) + 1024;

__device__ __forceinline__ void ds_cbar() { asm volatile("bar.sync 1, %0;" ::"n"(DS_CTHREADS) : "memory"); }
__device__ __forceinline__ void ds_cp_async_16(uint32_t dst_smem, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst_smem), "l"(src) : "memory");
}

// The scratch sits behind an integer round-up of the dynamic shared-memory base, so the compiler no longer knows its address
// space and would emit GENERIC loads / stores for it (LD.E / ST.E: measured 2 900 cycles per item for the 60 loads of the merge).
// Everything in the item loop therefore goes through explicit shared-space accesses.
__device__ __forceinline__ float ds_lds(uint32_t a) { float v; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a)); return v; }
__device__ __forceinline__ float2 ds_lds2(uint32_t a) {
  float2 v;
  asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a));
  return v;
}
__device__ __forceinline__ void ds_sts(uint32_t a, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory"); }
__device__ __forceinline__ void ds_sts2(uint32_t a, float x, float y) { asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(a), "f"(x), "f"(y) : "memory"); }

static_assert(DS_CWARPS == DS_STAGES, "one consumer warp per ring stage");
__device__ unsigned long long g_ds_trace[8];     // DH_DEC_ATTN_DBG=3: cycles of the compute warps by section
__device__ unsigned long long g_ds_span[2][512][4];   // DH_DEC_ATTN_DBG=3: globaltimer at block entry / loop begin / loop end of the last two launches
__device__ __forceinline__ unsigned long long ds_gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// 16 warps = 4 per SM sub-partition, whose 16384 registers then allow 128 each (a 17th warp would cap all of them at 96)
template <int DBG>
__global__ void __launch_bounds__(DS_THREADS, 1)
attn_decode_stream_kernel(const __grid_constant__ CUtensorMap kmap, const __grid_constant__ CUtensorMap vmap, const bf16* __restrict__ qkv,
                          const int* __restrict__ lens, bf16* __restrict__ out, const bf16* __restrict__ prefix_kv,
                          const float* __restrict__ gate, int nseq, int max_len, int G, int qpk, int prefix_len, long long ld_qkv, float scale, int dbg_slot) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* ring = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  DsScratch* sc = reinterpret_cast<DsScratch*>(ring + DS_STAGES * DS_STAGE_BYTES);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (DBG == 3 && tid == 0 && blockIdx.x < 512) g_ds_span[dbg_slot][blockIdx.x][0] = ds_gtime();
  // gridDim.x is a multiple of G: the block keeps ONE kv group (its adapter prefix rows are staged once) and walks the sequences
  // seq0, seq0 + seq_step, ...  The chunks of all its items form one stream n = 0, 1, ... through the ring (stage n % STAGES,
  // consumer warp n % CWARPS), so the producer runs ahead into the next item while the compute warps merge the current one.
  const int g = blockIdx.x % G, seq0 = blockIdx.x / G, seq_step = gridDim.x / G;
  if (tid == 0) {
    for (int i = 0; i < DS_STAGES; ++i) { mbar_init(&sc->full[i], 1); mbar_init(&sc->empty[i], 1); }
    fence_barrier_init();
  }
  __syncthreads();
  pdl_launch_dependents();
  // lens was written before the previous kernel started (the chain of pdl_wait is transitive)

  if (warp == DS_CWARPS) {
    // ===================== K / V chunk producer =====================
    if (lane == 0) {
      bool waited = false;
      int n = 0;
      for (int seq = seq0; seq < nseq; seq += seq_step) {
        const int nch = (__ldg(lens + seq) + DS_KEYS - 1) / DS_KEYS;
        const int row0 = (seq * G + g) * max_len;
        for (int c = 0; c < nch; ++c, ++n) {
          if (c == nch - 1 && !waited) { pdl_wait(); waited = true; }   // only an item's last chunk holds the key the previous kernel wrote
          const int s = n % DS_STAGES;
          mbar_wait(&sc->empty[s], ((n / DS_STAGES) & 1) ^ 1);
          if (DBG == 2) { mbar_arrive(&sc->full[s]); continue; }
          mbar_expect_tx(&sc->full[s], DS_STAGE_BYTES);
          tma_load_2d(ring + s * DS_STAGE_BYTES, &kmap, 0, row0 + c * DS_KEYS, &sc->full[s]);
          tma_load_2d(ring + s * DS_STAGE_BYTES + DS_KEYS * 128, &vmap, 0, row0 + c * DS_KEYS, &sc->full[s]);
        }
      }
      if (!waited) pdl_wait();
    }
    return;
  }
  // ===================== compute warps =====================
  const int r = lane >> 2, q = lane & 3;
  const float sl2 = scale * 1.4426950408889634f;
  const int H = G * qpk;
  const long long qoff = (long long)(g * (qpk + 2)) * 64;
  // the group's prefix K / V rows as one more chunk (16-byte pieces XOR-swizzled by row & 7 like the TMA boxes; rows past
  // prefix_len zero: a masked key has p = 0, and 0 * V must not meet a NaN)
  const uint32_t pfx_base = smem_u32(sc->pfx);
  const uint32_t part_a = smem_u32(sc->part), pm_a = smem_u32(sc->pm), pl_a = smem_u32(sc->pl), spo_a = smem_u32(sc->spo);
  constexpr uint32_t PART_W = DS_MAX_QPK * 64 * 4, PM_W = DS_MAX_QPK * 4;      // bytes per warp in part / pm / pl
  if (prefix_len > 0) {
    for (int i = tid; i < DS_KEYS * 16; i += DS_CTHREADS) {
      const int a = i >> 4, c = i & 15, kv = c >> 3, c8 = c & 7;       // row, 16-byte piece of K (kv = 0) or V (kv = 1)
      const uint32_t dst = pfx_base + kv * (DS_KEYS * 128) + a * 128 + ((c8 ^ (a & 7)) << 4);
      if (a < prefix_len) ds_cp_async_16(dst, prefix_kv + (long long)a * ld_qkv + qoff + (long long)(qpk + kv) * 64 + c8 * 8);
      else asm volatile("st.shared.v4.u32 [%0], {%1, %1, %1, %1};" ::"r"(dst), "r"(0u) : "memory");
    }
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  const float gate_r = (prefix_len > 0 && r < qpk) ? __ldg(gate + g * qpk + r) : 0.f;
  // everything above reads what was written before the token step began (the projected prefix, the gate): it runs under the tail of
  // the previous kernel; q is that kernel's output
  pdl_wait();
  uint32_t qa[4][2], qn[4][2];
  auto load_q = [&](uint32_t (&dst)[4][2], int seq) {
    const bf16* qrow = qkv + (long long)seq * ld_qkv + qoff;
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
      dst[ks][0] = r < qpk ? __ldcg(reinterpret_cast<const unsigned int*>(qrow + r * 64 + 16 * ks + 2 * q)) : 0u;
      dst[ks][1] = r < qpk ? __ldcg(reinterpret_cast<const unsigned int*>(qrow + r * 64 + 16 * ks + 8 + 2 * q)) : 0u;
    }
  };
  int len = 0, len_next = 0;
  if (seq0 < nseq) { load_q(qa, seq0); len = __ldg(lens + seq0); }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  ds_cbar();                                    // prefix chunk visible to every compute warp
  int n0 = 0;                                   // stream index of the item's first chunk
  long long t_wait = 0, t_chunk = 0, t_epi = 0, t_pfx = 0, n_chunks = 0, n_items = 0;
  const long long t_begin = clock64();
  if (DBG == 3 && tid == 0 && blockIdx.x < 512) g_ds_span[dbg_slot][blockIdx.x][1] = ds_gtime();
  for (int seq = seq0; seq < nseq; seq += seq_step) {
    const int nch = (len + DS_KEYS - 1) / DS_KEYS;
    const bool more = seq + seq_step < nseq;
    if (more) { load_q(qn, seq + seq_step); len_next = __ldg(lens + seq + seq_step); }     // in flight under this item's chunks
    if (DBG != 5 && prefix_len > 0 && warp == (n0 + nch) % DS_CWARPS) {
      // the warp next in the round-robin (it has the fewest chunks of this item) takes the prefix: own softmax, gate, no scale-in
      // with the cached keys (ger/adapter.py:116-146)
      KvcState sp;
      sp.init();
      kvc_chunk_step<DS_KEYS>(sp, qa, pfx_base, 0, prefix_len, sl2, lane);
      sp.l += __shfl_xor_sync(0xffffffffu, sp.l, 1);
      sp.l += __shfl_xor_sync(0xffffffffu, sp.l, 2);
      const float gs = gate_r / sp.l;
#pragma unroll
      for (int dt = 0; dt < 8; ++dt) ds_sts2(spo_a + (r * 64 + 8 * dt + 2 * q) * 4, gs * sp.acc[dt][0], gs * sp.acc[dt][1]);
    }
    KvcState st;
    st.init();
    // the item's chunks c with (n0 + c) % CWARPS == warp.  Stage (n0 + c) % STAGES == warp: a parity wait may look at most one
    // phase ahead, and with one consumer per stage the warp's own program order guarantees that
    for (int c = (warp - n0 % DS_CWARPS + DS_CWARPS) % DS_CWARPS; c < nch; c += DS_CWARPS) {
      const int n = n0 + c;
      const long long ta = DBG == 3 ? clock64() : 0;
      mbar_wait(&sc->full[warp], (n / DS_STAGES) & 1);
      const long long tb = DBG == 3 ? clock64() : 0;
      if (DBG != 1) kvc_chunk_step<DS_KEYS>(st, qa, smem_u32(ring + warp * DS_STAGE_BYTES), c * DS_KEYS, len, sl2, lane);
      __syncwarp();
      if (lane == 0) mbar_arrive(&sc->empty[warp]);
      if (DBG == 3) { t_wait += tb - ta; t_chunk += clock64() - tb; ++n_chunks; }
    }
    n0 += nch;
    const long long tc = DBG == 3 ? clock64() : 0;
    st.l += __shfl_xor_sync(0xffffffffu, st.l, 1);
    st.l += __shfl_xor_sync(0xffffffffu, st.l, 2);
#pragma unroll
    for (int dt = 0; dt < 8; ++dt) ds_sts2(part_a + warp * PART_W + (r * 64 + 8 * dt + 2 * q) * 4, st.acc[dt][0], st.acc[dt][1]);
    if (q == 0) { ds_sts(pm_a + warp * PM_W + r * 4, st.m); ds_sts(pl_a + warp * PM_W + r * 4, st.l); }
    ds_cbar();
    const long long td = DBG == 3 ? clock64() : 0;
    // merge of the warps' running states, plus the prefix term: two adjacent columns per thread (threads past qpk * 32 idle)
    float o0 = 0.f, o1 = 0.f;
    const int mh = tid >> 5, md = (tid & 31) * 2;
    if (DBG != 4 && tid < qpk * 32) {
      float M = -INFINITY;
#pragma unroll
      for (int w = 0; w < DS_CWARPS; ++w) M = fmaxf(M, ds_lds(pm_a + w * PM_W + mh * 4));
      float num0 = 0.f, num1 = 0.f, den = 0.f;
#pragma unroll
      for (int w = 0; w < DS_CWARPS; ++w) {
        const float e = exp2f(ds_lds(pm_a + w * PM_W + mh * 4) - M);         // a warp without chunks left m = -inf: weight 0
        const float2 pv = ds_lds2(part_a + w * PART_W + (mh * 64 + md) * 4);
        num0 += e * pv.x;
        num1 += e * pv.y;
        den += e * ds_lds(pl_a + w * PM_W + mh * 4);
      }
      const float inv = 1.f / den;
      o0 = num0 * inv;
      o1 = num1 * inv;
      if (prefix_len > 0) {
        const float2 po = ds_lds2(spo_a + (mh * 64 + md) * 4);
        o0 += po.x;
        o1 += po.y;
      }
    }
    ds_cbar();       // orders this item's reads of part / pm / pl / spo before the next item's writes (the store comes after it)
    if (tid < qpk * 32)
      *reinterpret_cast<__nv_bfloat162*>(out + (long long)seq * H * 64 + (g * qpk + mh) * 64 + md) = __floats2bfloat162_rn(o0, o1);
    if (more) {
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) { qa[ks][0] = qn[ks][0]; qa[ks][1] = qn[ks][1]; }
      len = len_next;
    }
    if (DBG == 3) { t_epi += td - tc; t_pfx += clock64() - td; ++n_items; }
  }
  if (DBG == 3 && tid == 0 && blockIdx.x < 512) g_ds_span[dbg_slot][blockIdx.x][2] = ds_gtime();
  if (DBG == 3 && lane == 0) {
    atomicAdd(&g_ds_trace[0], (unsigned long long)t_wait);
    atomicAdd(&g_ds_trace[1], (unsigned long long)t_chunk);
    atomicAdd(&g_ds_trace[2], (unsigned long long)t_epi);
    atomicAdd(&g_ds_trace[3], (unsigned long long)t_pfx);
    atomicAdd(&g_ds_trace[4], (unsigned long long)(clock64() - t_begin));
    atomicAdd(&g_ds_trace[5], (unsigned long long)n_chunks);
    atomicAdd(&g_ds_trace[6], (unsigned long long)n_items);
    atomicAdd(&g_ds_trace[7], 1ull);
  }
}

}  // namespace dh

using namespace dh;

// Called by dh_attention_decode (attention.cu) when the cache rows are a multiple of 64 keys long.  Returns 0 on success.
extern "C" int dh_attention_decode_stream(const void* qkv, const void* cache_k, const void* cache_v, const int32_t* lens, void* out,
                                          const void* prefix_kv, const float* gate, int nseq, int max_len, int n_groups, int q_per_kv,
                                          int prefix_len, void* stream) {
  DH_CHECK(qkv && cache_k && cache_v && lens && out && nseq > 0, "dh_attention_decode_stream: bad argument");
  DH_CHECK(q_per_kv >= 1 && q_per_kv <= DS_MAX_QPK, "dh_attention_decode_stream: q_per_kv %d not in [1, %d]", q_per_kv, DS_MAX_QPK);
  DH_CHECK(prefix_len >= 0 && prefix_len <= DS_MAX_PREFIX, "dh_attention_decode_stream: prefix_len %d > %d", prefix_len, DS_MAX_PREFIX);
  DH_CHECK(max_len > 0 && max_len % DS_KEYS == 0, "dh_attention_decode_stream: max_len %d must be a multiple of %d", max_len, DS_KEYS);
  DH_CHECK((((uintptr_t)qkv | (uintptr_t)cache_k | (uintptr_t)cache_v | (uintptr_t)prefix_kv) & 15) == 0,
           "dh_attention_decode_stream: operands must be 16-byte aligned");
  static bool attr = false;
  if (!attr) {
    DH_CUDA(cudaFuncSetAttribute(attn_decode_stream_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, DS_SMEM));
    DH_CUDA(cudaFuncSetAttribute(attn_decode_stream_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, DS_SMEM));
    DH_CUDA(cudaFuncSetAttribute(attn_decode_stream_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, DS_SMEM));
    DH_CUDA(cudaFuncSetAttribute(attn_decode_stream_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, DS_SMEM));
    DH_CUDA(cudaFuncSetAttribute(attn_decode_stream_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, DS_SMEM));
    DH_CUDA(cudaFuncSetAttribute(attn_decode_stream_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, DS_SMEM));
    attr = true;
  }
  CUtensorMap kmap, vmap;
  const long long rows = (long long)nseq * n_groups * max_len;
  if (make_kv_cache_map(&kmap, cache_k, rows, DS_KEYS) || make_kv_cache_map(&vmap, cache_v, rows, DS_KEYS)) return -3;
  const long long ld = (long long)n_groups * (q_per_kv + 2) * 64;
  const bf16* pkv = (const bf16*)(prefix_len > 0 ? prefix_kv : nullptr);
  const int plen = pkv ? prefix_len : 0;
  // one resident block per SM, each walking its share of the (sequence, kv group) items: no second wave, no per-item launch ramp.
  // DH_DEC_ATTN_PERSIST=0: one block per item (A/B testing)
  static int persist = -1;
  if (persist < 0) {
    const char* e = getenv("DH_DEC_ATTN_PERSIST");
    persist = (e && e[0] == '0') ? 0 : 1;
  }
  int seq_blocks = nseq;
  if (persist) {
    const int cap = sm_count() / n_groups;
    if (cap >= 1 && seq_blocks > cap) seq_blocks = cap;
  }
  const char* dbg_e = getenv("DH_DEC_ATTN_DBG");
  const int dbg = dbg_e ? atoi(dbg_e) : 0;
  static int dbg_calls = 0;
  const int dbg_slot = dbg == 3 ? (dbg_calls++ & 1) : 0;
  auto kern = dbg == 1 ? attn_decode_stream_kernel<1> : dbg == 2 ? attn_decode_stream_kernel<2> : dbg == 3 ? attn_decode_stream_kernel<3> : dbg == 4 ? attn_decode_stream_kernel<4> : dbg == 5 ? attn_decode_stream_kernel<5> : attn_decode_stream_kernel<0>;
  DH_CUDA(launch_pdl(kern, dim3(seq_blocks * n_groups), dim3(DS_THREADS), (size_t)DS_SMEM, (cudaStream_t)stream, kmap, vmap,
                     (const bf16*)qkv, (const int*)lens, (bf16*)out, pkv, gate, nseq, max_len, n_groups, q_per_kv, plen, ld,
                     1.0f / sqrtf(64.0f), dbg_slot));
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_debug_ds_trace(unsigned long long* out8, unsigned long long* span) {
  DH_CUDA(cudaMemcpyFromSymbol(out8, dh::g_ds_trace, 8 * sizeof(unsigned long long)));
  if (span) DH_CUDA(cudaMemcpyFromSymbol(span, dh::g_ds_span, sizeof(dh::g_ds_span)));
  return 0;
}
