// Stand-alone streamed decode attention (the default behind dh_attention_decode): block = (sequence, kv group).  A producer
// warp streams the cached K / V rows of the item as 32-key chunks through an 8-stage TMA ring (the chunks that hold only keys of
// earlier token steps are requested BEFORE griddepcontrol.wait, so they overlap the tail of the QKV kernel); eight compute warps
// take the 32-key chunks round-robin (warp w always reads ring stage w), each keeping a flash-style running state for the group's q heads on mma.sync tiles
// (decode_attn.cuh), and merge in shared memory; the adapter prefix keeps its own softmax and gate (ger/adapter.py:116-146).
// 86 KB of shared memory and 288 threads per block: two blocks per SM, i.e. sixteen compute warps and 128 KB of chunks in flight per SM.
#include <math.h>
#include <stdlib.h>

#include "decode_attn.cuh"
#include "../../include/dualhyp.h"

namespace dh {

constexpr int DS_CWARPS = 8;            // == DS_STAGES: every ring stage is consumed by ONE warp (see the chunk loop)
constexpr int DS_CTHREADS = DS_CWARPS * 32;
constexpr int DS_THREADS = DS_CTHREADS + 32;
constexpr int DS_STAGES = 8;
constexpr int DS_KEYS = 32;             // keys per chunk: K 4 KB + V 4 KB per stage.  A warp's chunk is one long dependent chain (ldmatrix ->
                                        // mma -> softmax -> mma, ~5 us per 64 keys), so latency is hidden by warps, not by deeper chunks: the first
                                        // build (3 warps x 64 keys, 3 blocks per SM = 9 compute warps per SM) measured 48 us per layer at 256 sequences
constexpr int DS_STAGE_BYTES = 2 * DS_KEYS * 128;
constexpr int DS_MAX_QPK = 8;
constexpr int DS_MAX_PREFIX = 16;

struct DsScratch {
  float part[DS_CWARPS][DS_MAX_QPK][64];
  float pm[DS_CWARPS][DS_MAX_QPK];
  float pl[DS_CWARPS][DS_MAX_QPK];
  float sout[DS_MAX_QPK][64];
  bf16 spfx[DS_MAX_PREFIX][128];
  uint64_t full[DS_STAGES], empty[DS_STAGES];
};
constexpr int DS_SMEM = DS_STAGES * DS_STAGE_BYTES + (int)sizeof(DsScratch) + 1024;

__device__ __forceinline__ void ds_cbar() { asm volatile("bar.sync 1, %0;" ::"n"(DS_CTHREADS) : "memory"); }
__device__ __forceinline__ void ds_cp_async_16(uint32_t dst_smem, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst_smem), "l"(src) : "memory");
}

static_assert(DS_CWARPS == DS_STAGES, "one consumer warp per ring stage");

__global__ void __launch_bounds__(DS_THREADS, 2)
attn_decode_stream_kernel(const __grid_constant__ CUtensorMap kmap, const __grid_constant__ CUtensorMap vmap, const bf16* __restrict__ qkv,
                          const int* __restrict__ lens, bf16* __restrict__ out, const bf16* __restrict__ prefix_kv,
                          const float* __restrict__ gate, int max_len, int G, int qpk, int prefix_len, long long ld_qkv, float scale) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* ring = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  DsScratch* sc = reinterpret_cast<DsScratch*>(ring + DS_STAGES * DS_STAGE_BYTES);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int seq = blockIdx.x, g = blockIdx.y;
  if (tid == 0) {
    for (int i = 0; i < DS_STAGES; ++i) { mbar_init(&sc->full[i], 1); mbar_init(&sc->empty[i], 1); }
    fence_barrier_init();
  }
  __syncthreads();
  pdl_launch_dependents();
  const int len = __ldg(lens + seq);             // written before the previous kernel started (the chain of pdl_wait is transitive)
  const int nch = (len + DS_KEYS - 1) / DS_KEYS;
  const int row0 = (seq * G + g) * max_len;

  if (warp == DS_CWARPS) {
    // ===================== K / V chunk producer =====================
    if (lane == 0) {
      bool waited = false;
      for (int c = 0; c < nch; ++c) {
        if (c == nch - 1 && !waited) { pdl_wait(); waited = true; }     // only the last chunk holds the key the previous kernel wrote
        const int s = c % DS_STAGES;
        mbar_wait(&sc->empty[s], ((c / DS_STAGES) & 1) ^ 1);
        mbar_expect_tx(&sc->full[s], DS_STAGE_BYTES);
        tma_load_2d(ring + s * DS_STAGE_BYTES, &kmap, 0, row0 + c * DS_KEYS, &sc->full[s]);
        tma_load_2d(ring + s * DS_STAGE_BYTES + DS_KEYS * 128, &vmap, 0, row0 + c * DS_KEYS, &sc->full[s]);
      }
      if (!waited) pdl_wait();
    }
    return;
  }
  // ===================== compute warps =====================
  pdl_wait();
  const int r = lane >> 2, q = lane & 3;
  const float sl2 = scale * 1.4426950408889634f;
  const bf16* qrow = qkv + (long long)seq * ld_qkv + (long long)(g * (qpk + 2)) * 64;
  for (int i = tid; i < prefix_len * 16; i += DS_CTHREADS) {
    const int a = i >> 4, c = i & 15;
    ds_cp_async_16(smem_u32(&sc->spfx[a][c * 8]), prefix_kv + (long long)a * ld_qkv + (long long)(g * (qpk + 2) + qpk) * 64 + c * 8);
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  uint32_t qa[4][2];
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) {
    qa[ks][0] = r < qpk ? __ldcg(reinterpret_cast<const unsigned int*>(qrow + r * 64 + 16 * ks + 2 * q)) : 0u;
    qa[ks][1] = r < qpk ? __ldcg(reinterpret_cast<const unsigned int*>(qrow + r * 64 + 16 * ks + 8 + 2 * q)) : 0u;
  }
  float q0 = 0.f, q1 = 0.f, gate_h = 0.f;
  KvcState st;
  st.init();
  for (int c = warp; c < nch; c += DS_CWARPS) {
    const int s = c % DS_STAGES;
    // stage s == warp: a parity wait may look at most one phase ahead, and with one consumer per stage the warp's own program
    // order guarantees that (chunk c - STAGES was this warp's previous chunk)
    mbar_wait(&sc->full[s], (c / DS_STAGES) & 1);
    kvc_chunk_step<DS_KEYS>(st, qa, smem_u32(ring + s * DS_STAGE_BYTES), c * DS_KEYS, len, sl2, lane);
    __syncwarp();
    if (lane == 0) mbar_arrive(&sc->empty[s]);
  }
  st.l += __shfl_xor_sync(0xffffffffu, st.l, 1);
  st.l += __shfl_xor_sync(0xffffffffu, st.l, 2);
#pragma unroll
  for (int dt = 0; dt < 8; ++dt) *reinterpret_cast<float2*>(&sc->part[warp][r][8 * dt + 2 * q]) = make_float2(st.acc[dt][0], st.acc[dt][1]);
  if (q == 0) { sc->pm[warp][r] = st.m; sc->pl[warp][r] = st.l; }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  ds_cbar();
  for (int i = tid; i < qpk * 64; i += DS_CTHREADS) {
    const int h = i >> 6, d = i & 63;
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < DS_CWARPS; ++w) M = fmaxf(M, sc->pm[w][h]);
    float num = 0.f, den = 0.f;
#pragma unroll
    for (int w = 0; w < DS_CWARPS; ++w) {
      const float e = exp2f(sc->pm[w][h] - M);          // a warp without chunks left m = -inf: weight 0
      num += e * sc->part[w][h][d];
      den += e * sc->pl[w][h];
    }
    sc->sout[h][d] = num / den;
  }
  ds_cbar();
  const int H = G * qpk;
  for (int hh = warp; hh < qpk; hh += DS_CWARPS) {
    float pfx0 = 0.f, pfx1 = 0.f;
    if (prefix_len > 0) {
      q0 = __bfloat162float(qrow[hh * 64 + lane]) * scale;
      q1 = __bfloat162float(qrow[hh * 64 + lane + 32]) * scale;
      gate_h = gate[g * qpk + hh];
      float ps[DS_MAX_PREFIX];
      float pmx = -INFINITY;
#pragma unroll
      for (int a = 0; a < DS_MAX_PREFIX; ++a) {
        ps[a] = -INFINITY;
        if (a < prefix_len) {
          ps[a] = warp_sum(q0 * __bfloat162float(sc->spfx[a][lane]) + q1 * __bfloat162float(sc->spfx[a][lane + 32]));
          pmx = fmaxf(pmx, ps[a]);
        }
      }
      float psum = 0.f;
#pragma unroll
      for (int a = 0; a < DS_MAX_PREFIX; ++a) {
        ps[a] = (a < prefix_len) ? __expf(ps[a] - pmx) : 0.f;
        psum += ps[a];
      }
      const float gs = gate_h / psum;
#pragma unroll
      for (int a = 0; a < DS_MAX_PREFIX; ++a) {
        if (a < prefix_len) {
          pfx0 += gs * ps[a] * __bfloat162float(sc->spfx[a][64 + lane]);
          pfx1 += gs * ps[a] * __bfloat162float(sc->spfx[a][96 + lane]);
        }
      }
    }
    bf16* orow = out + (long long)seq * H * 64 + (g * qpk + hh) * 64;
    orow[lane] = __float2bfloat16_rn(sc->sout[hh][lane] + pfx0);
    orow[lane + 32] = __float2bfloat16_rn(sc->sout[hh][lane + 32] + pfx1);
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
    DH_CUDA(cudaFuncSetAttribute(attn_decode_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DS_SMEM));
    attr = true;
  }
  CUtensorMap kmap, vmap;
  const long long rows = (long long)nseq * n_groups * max_len;
  if (make_kv_cache_map(&kmap, cache_k, rows, DS_KEYS) || make_kv_cache_map(&vmap, cache_v, rows, DS_KEYS)) return -3;
  const long long ld = (long long)n_groups * (q_per_kv + 2) * 64;
  const bf16* pkv = (const bf16*)(prefix_len > 0 ? prefix_kv : nullptr);
  const int plen = pkv ? prefix_len : 0;
  DH_CUDA(launch_pdl(attn_decode_stream_kernel, dim3(nseq, n_groups), dim3(DS_THREADS), (size_t)DS_SMEM, (cudaStream_t)stream, kmap, vmap,
                     (const bf16*)qkv, (const int*)lens, (bf16*)out, pkv, gate, max_len, n_groups, q_per_kv, plen, ld, 1.0f / sqrtf(64.0f)));
  DH_CUDA(cudaGetLastError());
  return 0;
}
