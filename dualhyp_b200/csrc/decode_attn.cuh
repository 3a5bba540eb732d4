// Shared pieces of the streamed decode attention (ger/model.py:248-261 with one new token per sequence): a warp consumes one
// chunk of 64 cached keys that TMA laid down in shared memory (K tile [64 keys][64] bf16 at `kbase`, V tile 8 KB above it;
// 128-byte rows whose 16-byte chunks are XOR-swizzled by row & 7, i.e. CU_TENSOR_MAP_SWIZZLE_128B) and folds it into a
// flash-style running (max, sum, output) of the kv group's q heads.  The heads are the rows of mma.sync m16n8k16 tiles, so one
// K / V row serves all heads of the group (native GQA).  Used by csrc/decode.cu (persistent token-step
// kernel) and csrc/decode_attn.cu (stand-alone kernel behind dh_attention_decode).
#pragma once
#include "common.cuh"

namespace dh {

constexpr int KVC_KEYS = 64;                      // keys per chunk of the persistent kernel (csrc/decode.cu)
constexpr int KVC_BYTES = 2 * KVC_KEYS * 128;     // K tile + V tile = one 16 KB ring stage

__device__ __forceinline__ void kvc_mma_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void kvc_ldmatrix_x4(uint32_t (&r)[4], uint32_t smem_addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_addr));
}
__device__ __forceinline__ void kvc_ldmatrix_x4_trans(uint32_t (&r)[4], uint32_t smem_addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_addr));
}

// Running state of one warp: row r = lane >> 2 is the head (rows >= q_per_kv are idle), columns 8 dt + 2 (lane & 3) + {0, 1}.
struct KvcState {
  float m, l;            // running max (log2 domain) and per-lane partial sum of exp2
  float acc[8][4];       // O fragments ([dt][0..1] valid rows, [2..3] the idle rows 8-15)
  __device__ __forceinline__ void init() {
    m = -INFINITY;
    l = 0.f;
#pragma unroll
    for (int dt = 0; dt < 8; ++dt)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[dt][j] = 0.f;
  }
};

// qa[ks] = the A fragments of the q rows for d = 16 ks .. 16 ks + 15 ({a0, a2}; a1 = a3 = 0 are the idle rows).
// key0 = index of the chunk's first key, n_valid = keys [0, n_valid) exist, sl2 = softmax scale * log2(e).
template <int KEYS = KVC_KEYS>
__device__ __forceinline__ void kvc_chunk_step(KvcState& st, const uint32_t (&qa)[4][2], uint32_t kbase, int key0, int n_valid, float sl2, int lane) {
  static_assert(KEYS % 32 == 0, "a chunk is a whole number of 32-key halves");
  const uint32_t vbase = kbase + KEYS * 128;
  const int q = lane & 3, lr = lane & 7, lm = lane >> 3;
  // halves of 32 keys, each an online-softmax step of its own (keeps the score fragments at 16 registers)
#pragma unroll
  for (int hf = 0; hf < KEYS / 32; ++hf) {
    if (key0 + 32 * hf >= n_valid) break;               // warp-uniform: no valid key in this half
    float sc4[4][4];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
#pragma unroll
      for (int j = 0; j < 4; ++j) sc4[nt][j] = 0.f;
      uint32_t b[4];
      const uint32_t rowaddr = kbase + (32 * hf + 8 * nt + lr) * 128;
      kvc_ldmatrix_x4(b, rowaddr + ((lm ^ lr) << 4));
      kvc_mma_16816(sc4[nt], qa[0][0], 0u, qa[0][1], 0u, b[0], b[1]);
      kvc_mma_16816(sc4[nt], qa[1][0], 0u, qa[1][1], 0u, b[2], b[3]);
      kvc_ldmatrix_x4(b, rowaddr + (((4 + lm) ^ lr) << 4));
      kvc_mma_16816(sc4[nt], qa[2][0], 0u, qa[2][1], 0u, b[0], b[1]);
      kvc_mma_16816(sc4[nt], qa[3][0], 0u, qa[3][1], 0u, b[2], b[3]);
    }
    const int kb = key0 + 32 * hf + 2 * q;
    float mx = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const float v = (kb + 8 * nt + j < n_valid) ? sc4[nt][j] * sl2 : -INFINITY;
        sc4[nt][j] = v;
        mx = fmaxf(mx, v);
      }
    }
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
    const float m_new = fmaxf(st.m, mx);                // finite: the half's first key is valid for every row
    const float alpha = exp2f(st.m - m_new);
    st.m = m_new;
    st.l *= alpha;
#pragma unroll
    for (int dt = 0; dt < 8; ++dt) { st.acc[dt][0] *= alpha; st.acc[dt][1] *= alpha; }
    uint32_t pa[2][2];
#pragma unroll
    for (int kk = 0; kk < 2; ++kk) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const float p0 = exp2f(sc4[2 * kk + h][0] - m_new), p1 = exp2f(sc4[2 * kk + h][1] - m_new);
        st.l += p0 + p1;
        pa[kk][h] = pack_bf16x2(p0, p1);
      }
    }
#pragma unroll
    for (int kk = 0; kk < 2; ++kk) {
      const uint32_t rowaddr = vbase + (32 * hf + 16 * kk + (lm & 1) * 8 + lr) * 128;
#pragma unroll
      for (int dp = 0; dp < 4; ++dp) {
        uint32_t b[4];
        kvc_ldmatrix_x4_trans(b, rowaddr + (((2 * dp + (lm >> 1)) ^ lr) << 4));
        kvc_mma_16816(st.acc[2 * dp], pa[kk][0], 0u, pa[kk][1], 0u, b[0], b[1]);
        kvc_mma_16816(st.acc[2 * dp + 1], pa[kk][0], 0u, pa[kk][1], 0u, b[2], b[3]);
      }
    }
  }
}

// KV cache [rows = n_seq * n_groups * max_len, 64] bf16 as rows of 128 bytes; box = 64 keys, 128-byte swizzle
inline int make_kv_cache_map(CUtensorMap* map, const void* base, long long rows, int box_keys = KVC_KEYS) {
  uint64_t dims[2] = {64, (uint64_t)rows};
  uint64_t strides[1] = {128};
  uint32_t box[2] = {64, (uint32_t)box_keys};
  return make_tensor_map(map, base, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

}  // namespace dh
