// Bandwidth-bound leaf kernels of the decoder path: RMSNorm fwd/bwd, embedding gather, RoPE (standalone
// API and in-place on the interleaved QKV row), cross-entropy over bf16 logits with in-place gradient,
// loss reduction, arg-max, casts, fused AdamW, KV-cache scatter.  All are one pass over their operands
// with 16-byte vector accesses; grids are sized by rows so every SM streams.
#include <math.h>

#include "common.cuh"
#include "../../include/dualhyp.h"

namespace dh {

// block-wide sum; every thread gets the result.  blockDim.x <= 1024, multiple of 32.
__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  float t = (l < nw) ? red[l] : 0.f;
  return warp_sum(t);
}
__device__ __forceinline__ float block_max(float v, float* red) {
  v = warp_max(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  float t = (l < nw) ? red[l] : -INFINITY;
  return warp_max(t);
}

__device__ __forceinline__ void load8(const float* p, float (&f)[8]) {
  float4 a = reinterpret_cast<const float4*>(p)[0], b = reinterpret_cast<const float4*>(p)[1];
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
__device__ __forceinline__ void load8(const bf16* p, float (&f)[8]) {
  uint4 u = *reinterpret_cast<const uint4*>(p);
  float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ void store8(float* p, const float (&f)[8]) {
  reinterpret_cast<float4*>(p)[0] = make_float4(f[0], f[1], f[2], f[3]);
  reinterpret_cast<float4*>(p)[1] = make_float4(f[4], f[5], f[6], f[7]);
}
__device__ __forceinline__ void store8(bf16* p, const float (&f)[8]) {
  uint4 u;
  u.x = pack_bf16x2(f[0], f[1]); u.y = pack_bf16x2(f[2], f[3]);
  u.z = pack_bf16x2(f[4], f[5]); u.w = pack_bf16x2(f[6], f[7]);
  *reinterpret_cast<uint4*>(p) = u;
}

// ---------------------------------------------------------------------------------------------
// RMSNorm (ger/rmsnorm.py:17-21).  One block per row, 8 elements per thread (blockDim = C/8).
// ---------------------------------------------------------------------------------------------
template <typename XT>
__global__ void rmsnorm_fwd_kernel(const XT* __restrict__ x, const float* __restrict__ w, bf16* __restrict__ y,
                                   float* __restrict__ rstd, int C, float eps) {
  __shared__ float red[32];
  const long long row = blockIdx.x;
  const int c0 = threadIdx.x * 8;
  float v[8], g[8];
  pdl_launch_dependents();
  pdl_wait();
  load8(x + row * C + c0, v);
  load8(w + c0, g);
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) ss += v[i] * v[i];
  ss = block_sum(ss, red);
  const float r = rsqrtf(ss / (float)C + eps);
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = g[i] * (v[i] * r);
  store8(y + row * C + c0, v);
  if (rstd && threadIdx.x == 0) rstd[row] = r;
}

// Many rows (training step, prefill): one WARP per row, C = NCH * 256, lane l owns the 8-element chunks j * 32 + l.  All NCH
// 32-byte loads of a lane are in flight before the first add and the row sum is five shuffles (no block barrier): the
// one-block-per-row form above keeps a single load per thread in flight and measured 3.7 TB/s on [8192, 2048] fp32.
template <typename XT, int NCH>
__global__ void __launch_bounds__(256) rmsnorm_fwd_warp_kernel(const XT* __restrict__ x, const float* __restrict__ w, bf16* __restrict__ y,
                                                               float* __restrict__ rstd, int M, float eps) {
  constexpr int C = NCH * 256;
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const long long row = blockIdx.x * 8LL + (threadIdx.x >> 5);
  if (row >= M) return;
  const XT* xr = x + row * C;
  float v[NCH][8];
#pragma unroll
  for (int j = 0; j < NCH; ++j) load8(xr + (j * 32 + lane) * 8, v[j]);
  float ss = 0.f;
#pragma unroll
  for (int j = 0; j < NCH; ++j)
#pragma unroll
    for (int i = 0; i < 8; ++i) ss += v[j][i] * v[j][i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
  const float r = rsqrtf(ss / (float)C + eps);
  bf16* yr = y + row * C;
#pragma unroll
  for (int j = 0; j < NCH; ++j) {
    float g[8];
    load8(w + (j * 32 + lane) * 8, g);
#pragma unroll
    for (int i = 0; i < 8; ++i) v[j][i] = g[i] * (v[j][i] * r);
    store8(yr + (j * 32 + lane) * 8, v[j]);
  }
  if (rstd && lane == 0) rstd[row] = r;
}

template <typename XT>
static cudaError_t launch_rmsnorm_fwd(const XT* x, const float* w, bf16* y, float* rstd, int M, int C, float eps, cudaStream_t s) {
  const dim3 grid((M + 7) / 8), block(256);
  if (M >= 512) {
    if (C == 2048) return launch_pdl(rmsnorm_fwd_warp_kernel<XT, 8>, grid, block, 0, s, x, w, y, rstd, M, eps);
    if (C == 1024) return launch_pdl(rmsnorm_fwd_warp_kernel<XT, 4>, grid, block, 0, s, x, w, y, rstd, M, eps);
    if (C == 512) return launch_pdl(rmsnorm_fwd_warp_kernel<XT, 2>, grid, block, 0, s, x, w, y, rstd, M, eps);
    if (C == 256) return launch_pdl(rmsnorm_fwd_warp_kernel<XT, 1>, grid, block, 0, s, x, w, y, rstd, M, eps);
  }
  return launch_pdl(rmsnorm_fwd_kernel<XT>, dim3(M), dim3(C / 8), 0, s, x, w, y, rstd, C, eps);
}

// Decode: consumer of a split-K GEMM (dh_gemm with split_k > 1 leaves `splits` fp32 partial products [splits][M][C]):
//   x_out = resid + bf16(sum_s partial[s])      (the residual epilogue of the GEMM, ger/model.py:185-186)
//   y     = rmsnorm(x_out) * w  in bf16         (the next norm, ger/rmsnorm.py:17-21)
// One row per block; the partials are summed in split order (deterministic).
__global__ void splitk_resid_rmsnorm_kernel(const float* __restrict__ partial, int splits, long long split_stride,
                                            const float* __restrict__ resid, const float* __restrict__ w, float* __restrict__ x_out,
                                            bf16* __restrict__ y, int C, float eps) {
  __shared__ float red[32];
  pdl_launch_dependents();
  pdl_wait();
  const long long row = blockIdx.x;
  const int c0 = threadIdx.x * 8;
  float acc[8], v[8], g[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  const float* prow = partial + row * C + c0;
  for (int s0 = 0; s0 < splits; s0 += 8) {          // 8 splits (16 independent 16-byte loads) in flight per thread
    float t[8][8];
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (s0 + u < splits) {
        const float4 a = __ldcg(reinterpret_cast<const float4*>(prow + (s0 + u) * split_stride));
        const float4 b = __ldcg(reinterpret_cast<const float4*>(prow + (s0 + u) * split_stride) + 1);
        t[u][0] = a.x; t[u][1] = a.y; t[u][2] = a.z; t[u][3] = a.w; t[u][4] = b.x; t[u][5] = b.y; t[u][6] = b.z; t[u][7] = b.w;
      }
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (s0 + u < splits) {
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] += t[u][i];
      }
  }
  load8(resid + row * C + c0, v);
  load8(w + c0, g);
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    v[i] += __bfloat162float(__float2bfloat16_rn(acc[i]));
    ss += v[i] * v[i];
  }
  store8(x_out + row * C + c0, v);
  ss = block_sum(ss, red);
  const float r = rsqrtf(ss / (float)C + eps);
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = g[i] * (v[i] * r);
  store8(y + row * C + c0, v);
}

// Decode: consumer of the split-K QKV GEMM.  One block per token row, one thread per 8 columns of a 64-wide head slot:
//   qkv = bf16(sum_s partial[s])  -> NeoX RoPE on the q and k slots at the row's position (the DH_EPI_QKV_ROPE math:
//   bf16-rounded product, bf16 tables, pairs (j, j + 32)) -> bf16 qkv row, and the k / v slots also go to the KV cache at
//   (seq[row], pos[row]) -- reduction + ger/model.py:240-246 + the cache update :248-259 in one launch.
__global__ void splitk_qkv_rope_cache_kernel(const float* __restrict__ partial, int splits, long long split_stride, bf16* __restrict__ qkv,
                                             bf16* ck, bf16* cv, const int* __restrict__ seq, const int* __restrict__ pos,
                                             const bf16* __restrict__ rope_cos, const bf16* __restrict__ rope_sin, int N, int max_len,
                                             int n_groups, int qpk) {
  pdl_launch_dependents();
  pdl_wait();
  const long long row = blockIdx.x;
  const int p = pos[row];
  for (int t = threadIdx.x; t < N / 16; t += blockDim.x) {       // thread: columns [j, j + 8) and [j + 32, j + 40) of one head slot
    const int slot = t >> 2, j = (t & 3) * 8;
    const float* pr = partial + row * N + slot * 64 + j;
    float lo[8], hi[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) lo[i] = hi[i] = 0.f;
    for (int s0 = 0; s0 < splits; s0 += 4) {
      float4 a[4][4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (s0 + u < splits) {
          const float4* q4 = reinterpret_cast<const float4*>(pr + (s0 + u) * split_stride);
          a[u][0] = __ldcg(q4); a[u][1] = __ldcg(q4 + 1); a[u][2] = __ldcg(q4 + 8); a[u][3] = __ldcg(q4 + 9);
        }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (s0 + u < splits) {
          lo[0] += a[u][0].x; lo[1] += a[u][0].y; lo[2] += a[u][0].z; lo[3] += a[u][0].w;
          lo[4] += a[u][1].x; lo[5] += a[u][1].y; lo[6] += a[u][1].z; lo[7] += a[u][1].w;
          hi[0] += a[u][2].x; hi[1] += a[u][2].y; hi[2] += a[u][2].z; hi[3] += a[u][2].w;
          hi[4] += a[u][3].x; hi[5] += a[u][3].y; hi[6] += a[u][3].z; hi[7] += a[u][3].w;
        }
    }
    const int in_group = slot % (qpk + 2), g = slot / (qpk + 2);
    if (in_group <= qpk) {                                        // q heads and the k head rotate; v passes through
      float c[8], sn[8];
      load8(rope_cos + (long long)p * 64 + j, c);                 // table halves are identical (repeat(1, 2))
      load8(rope_sin + (long long)p * 64 + j, sn);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float x1 = __bfloat162float(__float2bfloat16_rn(lo[i])), x2 = __bfloat162float(__float2bfloat16_rn(hi[i]));
        lo[i] = x1 * c[i] - x2 * sn[i];
        hi[i] = x2 * c[i] + x1 * sn[i];
      }
    }
    bf16* dst = qkv + row * N + slot * 64 + j;
    store8(dst, lo);
    store8(dst + 32, hi);
    if (in_group >= qpk) {                                        // k or v slot: also into the cache
      bf16* cd = (in_group == qpk ? ck : cv) + (((long long)seq[row] * n_groups + g) * max_len + p) * 64 + j;
      store8(cd, lo);
      store8(cd + 32, hi);
    }
  }
}

// dx = dres + rstd * (w*dy - xhat * mean(w*dy*xhat)),  xhat = x*rstd
template <typename XT>
__global__ void rmsnorm_bwd_kernel(const XT* __restrict__ x, const float* __restrict__ w, const float* __restrict__ rstd,
                                   const bf16* __restrict__ dy, const float* dres, float* dx_out, bf16* dx_bf16, int C) {
  __shared__ float red[32];
  const long long row = blockIdx.x;
  const int c0 = threadIdx.x * 8;
  float xv[8], g[8], d[8];
  pdl_launch_dependents();
  pdl_wait();
  load8(x + row * C + c0, xv);
  load8(w + c0, g);
  load8(dy + row * C + c0, d);
  const float r = rstd[row];
  float dot = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    xv[i] *= r;          // xhat
    d[i] *= g[i];        // w*dy
    dot += d[i] * xv[i];
  }
  dot = block_sum(dot, red) / (float)C;
  float o[8];
  if (dres) load8(dres + row * C + c0, o);
  else {
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i] = 0.f;
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) o[i] += r * (d[i] - xv[i] * dot);
  store8(dx_out + row * C + c0, o);
  if (dx_bf16) store8(dx_bf16 + row * C + c0, o);
}

// ---------------------------------------------------------------------------------------------
// embedding gather (ger/lora.py:537)
// ---------------------------------------------------------------------------------------------
template <typename WT, typename OT>
__global__ void embedding_kernel(const long long* __restrict__ idx, const WT* __restrict__ wte, OT* __restrict__ out, int C,
                                 int vocab_rows) {
  const long long row = blockIdx.x;
  long long id = idx[row];
  if (id < 0 || id >= vocab_rows) __trap();   // nn.Embedding raises on out-of-range ids
  for (int c0 = threadIdx.x * 8; c0 < C; c0 += blockDim.x * 8) {
    float v[8];
    load8(wte + id * C + c0, v);
    store8(out + row * C + c0, v);
  }
}

// ---------------------------------------------------------------------------------------------
// RoPE
// ---------------------------------------------------------------------------------------------
// apply_rotary_emb_func (ger/fused_rotary_embedding.py:10-90): x (B,T,H,D), cos/sin [T, rot/2].
template <typename CT>
__global__ void rope_apply_kernel(const bf16* __restrict__ x, bf16* out, const CT* __restrict__ cs, const CT* __restrict__ sn,
                                  long long total_pairs, int T, int H, int D, int rot, int interleaved, int conj) {
  const int half = rot / 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total_pairs; i += (long long)gridDim.x * blockDim.x) {
    const int p = (int)(i % half);
    const long long bth = i / half;
    const int t = (int)((bth / H) % T);
    const long long base = bth * D;
    const int i1 = interleaved ? 2 * p : p, i2 = interleaved ? 2 * p + 1 : p + half;
    const float c = (float)cs[(long long)t * half + p];
    float s = (float)sn[(long long)t * half + p];
    if (conj) s = -s;
    const float x1 = __bfloat162float(x[base + i1]), x2 = __bfloat162float(x[base + i2]);
    out[base + i1] = __float2bfloat16_rn(x1 * c - x2 * s);
    out[base + i2] = __float2bfloat16_rn(x1 * s + x2 * c);
  }
}
__global__ void copy_tail_kernel(const bf16* __restrict__ x, bf16* out, long long n_heads, int D, int rot) {
  const int tail = D - rot;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_heads * tail; i += (long long)gridDim.x * blockDim.x) {
    const long long h = i / tail;
    const int d = rot + (int)(i % tail);
    out[h * D + d] = x[h * D + d];
  }
}

// In-place NeoX rotation of the q and k slots of group-interleaved rows; tables bf16 [block, 64].
__global__ void rope_qkv_inplace_kernel(bf16* qkv, const bf16* __restrict__ cs, const bf16* __restrict__ sn,
                                        const int* __restrict__ positions, int M, int seq_len, int n_groups, int qpk, int conj) {
  const int slots = n_groups * (qpk + 2);
  const long long total = (long long)M * slots * 4;   // 4 threads per 64-wide head, 8 pairs each
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int part = (int)(i & 3);
    const long long rs = i >> 2;
    const int slot = (int)(rs % slots);
    const long long row = rs / slots;
    if ((slot % (qpk + 2)) == qpk + 1) continue;   // v slot
    const int pos = positions ? positions[row] : (int)(row % seq_len);
    bf16* h = qkv + (row * slots + slot) * 64;
    float a[8], b[8], c[8], s[8];
    load8(h + part * 8, a);
    load8(h + 32 + part * 8, b);
    load8(cs + (long long)pos * 64 + part * 8, c);
    load8(sn + (long long)pos * 64 + part * 8, s);
    float o1[8], o2[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float sj = conj ? -s[j] : s[j];
      o1[j] = a[j] * c[j] - b[j] * sj;
      o2[j] = b[j] * c[j] + a[j] * sj;
    }
    store8(h + part * 8, o1);
    store8(h + 32 + part * 8, o2);
  }
}

// ---------------------------------------------------------------------------------------------
// cross entropy over one bf16 logits row held in shared memory; gradient written in place
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) cross_entropy_kernel(bf16* logits, long long ldl, const long long* __restrict__ targets,
                                                            float* __restrict__ row_loss, int V, float grad_scale, int write_grad) {
  extern __shared__ uint4 srow[];   // V/8 uint4
  __shared__ float red[32];
  const long long row = blockIdx.x;
  const long long tg = targets[row];
  bf16* lrow = logits + row * ldl;
  const int nv = V / 8;
  if (tg < 0) {   // ignore_index = -1: no loss, zero gradient
    if (threadIdx.x == 0) row_loss[row] = 0.f;
    if (write_grad == 1) {
      const uint4 z = make_uint4(0, 0, 0, 0);
      for (int i = threadIdx.x; i < nv; i += blockDim.x) reinterpret_cast<uint4*>(lrow)[i] = z;
    }
    return;
  }
  if (tg >= V) __trap();
  float mx = -INFINITY;
  for (int i = threadIdx.x; i < nv; i += blockDim.x) {
    uint4 u = reinterpret_cast<const uint4*>(lrow)[i];
    srow[i] = u;
    float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
    mx = fmaxf(mx, fmaxf(fmaxf(fmaxf(a.x, a.y), fmaxf(b.x, b.y)), fmaxf(fmaxf(c.x, c.y), fmaxf(d.x, d.y))));
  }
  mx = block_max(mx, red);
  const float L2E = 1.4426950408889634f;
  float sum = 0.f;
  for (int i = threadIdx.x; i < nv; i += blockDim.x) {
    uint4 u = srow[i];
    float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
    sum += exp2f((a.x - mx) * L2E) + exp2f((a.y - mx) * L2E) + exp2f((b.x - mx) * L2E) + exp2f((b.y - mx) * L2E) +
           exp2f((c.x - mx) * L2E) + exp2f((c.y - mx) * L2E) + exp2f((d.x - mx) * L2E) + exp2f((d.y - mx) * L2E);
  }
  sum = block_sum(sum, red);
  const float lse = mx + logf(sum);
  const bf16* s16 = reinterpret_cast<const bf16*>(srow);
  if (threadIdx.x == 0) row_loss[row] = lse - __bfloat162float(s16[tg]);
  if (write_grad) {
    const float inv = grad_scale / sum;
    for (int i = threadIdx.x; i < nv; i += blockDim.x) {
      uint4 u = srow[i];
      float f[8];
      float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
      f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        f[j] = exp2f((f[j] - mx) * L2E) * inv;
        if ((long long)i * 8 + j == tg) f[j] -= grad_scale;
      }
      store8(lrow + (long long)i * 8, f);
    }
  }
}

__global__ void loss_reduce_kernel(const float* __restrict__ row_loss, const long long* __restrict__ targets, float* out2, int M) {
  __shared__ float red[32];
  float s = 0.f, n = 0.f;
  for (int i = threadIdx.x; i < M; i += blockDim.x) {
    s += row_loss[i];
    n += targets[i] >= 0 ? 1.f : 0.f;
  }
  s = block_sum(s, red);
  n = block_sum(n, red);
  if (threadIdx.x == 0) { out2[0] = s; out2[1] = n; }
}

__global__ void __launch_bounds__(256) argmax_kernel(const bf16* __restrict__ logits, long long ldl, long long* __restrict__ ids, int V) {
  __shared__ float sv[8];
  __shared__ int si[8];
  pdl_launch_dependents();
  pdl_wait();
  const long long row = blockIdx.x;
  const bf16* l = logits + row * ldl;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  if ((((uintptr_t)l) & 15) == 0 && (V & 7) == 0) {
    // 16-byte loads (8 logits), two in flight per thread; ascending order + strict '>' keep the lowest index per thread
    const uint4* l4 = reinterpret_cast<const uint4*>(l);
    const int n4 = V >> 3;
    for (int i = threadIdx.x; i < n4; i += 2 * blockDim.x) {
      const int i2 = i + blockDim.x;
      const uint4 a = __ldg(l4 + i);
      const uint4 b = i2 < n4 ? __ldg(l4 + i2) : make_uint4(0xff80ff80u, 0xff80ff80u, 0xff80ff80u, 0xff80ff80u);   // -inf
      float f[8];
      unpack8(a, f);
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (f[j] > best) { best = f[j]; bi = 8 * i + j; }
      unpack8(b, f);
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (f[j] > best) { best = f[j]; bi = 8 * i2 + j; }
    }
  } else {
    for (int i = threadIdx.x; i < V; i += blockDim.x) {
      float v = __bfloat162float(l[i]);
      if (v > best) { best = v; bi = i; }   // strided scan keeps the lowest index per thread
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float ov = __shfl_xor_sync(0xffffffffu, best, o);
    int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = bi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w)
      if (sv[w] > best || (sv[w] == best && si[w] < bi)) { best = sv[w]; bi = si[w]; }
    ids[row] = bi;
  }
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, long long lds, bf16* __restrict__ dst, long long ldd,
                                     long long rows, long long cols) {
  const long long cv = cols / 8;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < rows * cv; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cv, c = (i % cv) * 8;
    float v[8];
    load8(src + r * lds + c, v);
    store8(dst + r * ldd + c, v);
  }
}

// Philox4x32-10 (Salmon et al., SC'11): counter-based, so the mask of element block i is a pure function of (seed, offset, i) --
// the backward regenerates it instead of saving it.
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, ctr.x), lo0 = 0xD2511F53u * ctr.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, ctr.z), lo1 = 0xCD9E8D57u * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += 0x9E3779B9u;
    key.y += 0xBB67AE85u;
  }
  return ctr;
}

// nn.Dropout(p) on a bf16 tensor (ger/lora.py:94-98,165: the LoRA branch sees dropout(x)): out = keep ? x / (1 - p) : 0, one
// Philox call per 8 elements (eight 16-bit uniforms; keep when u16 >= p * 65536).  ADD = true: out += the masked, scaled x (the
// gradient of the LoRA branch flowing back into x through the same mask).
template <bool ADD>
__global__ void dropout_bf16_kernel(const bf16* __restrict__ x, bf16* __restrict__ out, long long n8, uint32_t thresh, float inv_keep,
                                    uint2 key, uint2 offset) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    const uint4 rnd = philox4x32_10(make_uint4((uint32_t)i, (uint32_t)(i >> 32), offset.x, offset.y), key);
    const uint32_t w[4] = {rnd.x, rnd.y, rnd.z, rnd.w};
    float v[8];
    unpack8(*reinterpret_cast<const uint4*>(x + i * 8), v);
    float o[8];
    if (ADD) unpack8(*reinterpret_cast<const uint4*>(out + i * 8), o);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const uint32_t u = (w[j >> 1] >> (16 * (j & 1))) & 0xFFFFu;
      const float m = u >= thresh ? v[j] * inv_keep : 0.f;
      o[j] = ADD ? o[j] + m : m;
    }
    *reinterpret_cast<uint4*>(out + i * 8) = pack8(o);
  }
}

// torch.optim.AdamW (decoupled weight decay, bias-corrected), one flat bucket
__global__ void adamw_kernel(float* p, const float* __restrict__ g, float* m, float* v, long long n, float lr, float b1, float b2,
                             float eps, float wd, float bc1, float bc2_sqrt, float gscale) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float gi = g[i] * gscale;
    float pi = p[i] * (1.f - lr * wd);
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] = pi - (lr / bc1) * (mi / denom);
  }
}

// scatter k and v slots of interleaved qkv rows into [nseq, n_groups, max_len, 64] caches
__global__ void kv_cache_write_kernel(const bf16* __restrict__ qkv, bf16* ck, bf16* cv, const int* __restrict__ seq,
                                      const int* __restrict__ pos, int M, int max_len, int n_groups, int qpk) {
  const long long total = (long long)M * n_groups * 2 * 8;   // 8 uint4 per 64-wide head
  const int slots = n_groups * (qpk + 2);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int part = (int)(i & 7);
    long long r = i >> 3;
    const int is_v = (int)(r & 1); r >>= 1;
    const int g = (int)(r % n_groups);
    const long long row = r / n_groups;
    const uint4 u = *reinterpret_cast<const uint4*>(qkv + (row * slots + g * (qpk + 2) + qpk + is_v) * 64 + part * 8);
    bf16* dst = (is_v ? cv : ck) + (((long long)seq[row] * n_groups + g) * max_len + pos[row]) * 64 + part * 8;
    *reinterpret_cast<uint4*>(dst) = u;
  }
}

static inline int grid_for(long long work, int block) {
  long long g = (work + block - 1) / block;
  long long cap = (long long)sm_count() * 16;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

// -------------------------------------------------------------------------------------------------
// Adapter prefix K/V projection for ALL adapter layers in one launch (ger/adapter.py:129-137: `self.attn(prefix)`
// then only the k and v slots are used).  wkv [L, G*128, C] holds the k and v rows of every layer's QKV weight
// (row r of group g: k rows 0..63, v rows 64..127).  HBM-bound on the weights: one warp per weight row, the aT
// prefix rows of the layer staged in shared memory.
// -------------------------------------------------------------------------------------------------
constexpr int PFX_MAX = 16;

__global__ void __launch_bounds__(256) prefix_project_fwd_kernel(const bf16* __restrict__ prefix, const bf16* __restrict__ wkv,
                                                                 bf16* __restrict__ out, int aT, int C, int R, int G, int qpk,
                                                                 long long ld_out, int rows_per_block) {
  extern __shared__ uint4 spfx4[];                    // [aT][C] bf16
  const int l = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint4* src = reinterpret_cast<const uint4*>(prefix + (long long)l * aT * C);
  for (int i = threadIdx.x; i < aT * C / 8; i += blockDim.x) spfx4[i] = src[i];
  __syncthreads();
  const int r0 = blockIdx.x * rows_per_block;
  for (int r = r0 + warp; r < r0 + rows_per_block && r < R; r += 8) {
    const uint4* wrow = reinterpret_cast<const uint4*>(wkv + ((long long)l * R + r) * C);
    float acc[PFX_MAX];
#pragma unroll
    for (int a = 0; a < PFX_MAX; ++a) acc[a] = 0.f;
    for (int c8 = lane; c8 < C / 8; c8 += 32) {
      float w[8];
      unpack8(__ldg(wrow + c8), w);
#pragma unroll
      for (int a = 0; a < PFX_MAX; ++a) {
        if (a < aT) {
          float x[8];
          unpack8(spfx4[a * (C / 8) + c8], x);
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[a] += w[i] * x[i];
        }
      }
    }
    const int g = r / 128, rr = r % 128;
    const long long col = (long long)(g * (qpk + 2) + qpk) * 64 + rr;
#pragma unroll
    for (int a = 0; a < PFX_MAX; ++a) {
      if (a < aT) {
        const float v = warp_sum(acc[a]);
        if (lane == 0) out[((long long)l * aT + a) * ld_out + col] = __float2bfloat16_rn(v);
      }
    }
  }
}

// d(prefix)[l, a, c] = sum_r d(prefix_kv)[l, a, col(r)] * wkv[l, r, c]   (fp32 out; thread = 2 columns c)
__global__ void __launch_bounds__(128) prefix_project_bwd_kernel(const float* __restrict__ dpkv, const bf16* __restrict__ wkv,
                                                                 float* __restrict__ dprefix, int aT, int C, int R, int G, int qpk,
                                                                 long long ld) {
  extern __shared__ float sd[];                       // [aT][R]
  const int l = blockIdx.y;
  for (int i = threadIdx.x; i < aT * R; i += blockDim.x) {
    const int a = i / R, r = i % R;
    const int g = r / 128, rr = r % 128;
    sd[i] = dpkv[((long long)l * aT + a) * ld + (long long)(g * (qpk + 2) + qpk) * 64 + rr];
  }
  __syncthreads();
  const int c = (blockIdx.x * 128 + threadIdx.x) * 2;
  if (c >= C) return;
  float a0[PFX_MAX], a1[PFX_MAX];
#pragma unroll
  for (int a = 0; a < PFX_MAX; ++a) { a0[a] = 0.f; a1[a] = 0.f; }
  const bf16* wl = wkv + (long long)l * R * C + c;
#pragma unroll 4
  for (int r = 0; r < R; ++r) {
    const float2 w = unpack_bf16x2(__ldg(reinterpret_cast<const uint32_t*>(wl + (long long)r * C)));
#pragma unroll
    for (int a = 0; a < PFX_MAX; ++a) {
      if (a < aT) {
        const float d = sd[a * R + r];
        a0[a] += d * w.x;
        a1[a] += d * w.y;
      }
    }
  }
#pragma unroll
  for (int a = 0; a < PFX_MAX; ++a)
    if (a < aT) *reinterpret_cast<float2*>(dprefix + ((long long)l * aT + a) * C + c) = make_float2(a0[a], a1[a]);
}

// -------------------------------------------------------------------------------------------------
// Valid-row compaction for the fused lm_head + loss: only rows whose target is not -1 contribute to the loss or to
// d(hidden), so only those are sent through the vocabulary projection.  One block: ordered compaction of the row
// indices (rows[i] = original row of the i-th valid target, -1 beyond the count), the gathered targets and the count.
// -------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) compact_targets_kernel(const long long* __restrict__ targets, int M, int* __restrict__ count,
                                                              int* __restrict__ rows, long long* __restrict__ tg_c) {
  __shared__ int warp_tot[32];
  __shared__ int base;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) base = 0;
  __syncthreads();
  for (int m0 = 0; m0 < M; m0 += 1024) {
    const int m = m0 + threadIdx.x;
    const long long tg = m < M ? targets[m] : -1;
    const bool ok = tg >= 0;
    const unsigned bal = __ballot_sync(0xffffffffu, ok);
    if (lane == 0) warp_tot[warp] = __popc(bal);
    __syncthreads();
    int off = base;
    for (int w = 0; w < warp; ++w) off += warp_tot[w];
    if (ok) {
      const int pos = off + __popc(bal & ((1u << lane) - 1));
      rows[pos] = m;
      tg_c[pos] = tg;
    }
    __syncthreads();
    if (threadIdx.x == 0) { int t = 0; for (int w = 0; w < 32; ++w) t += warp_tot[w]; base += t; }
    __syncthreads();
  }
  const int n = base;
  if (threadIdx.x == 0) *count = n;
  for (int i = n + threadIdx.x; i < M; i += 1024) { rows[i] = -1; tg_c[i] = -1; }
}

// dst[i, :] = src[rows[i], :] for i < *count  (bf16 rows of C elements, C % 8 == 0)
__global__ void __launch_bounds__(256) gather_rows_kernel(const bf16* __restrict__ src, const int* __restrict__ rows,
                                                          const int* __restrict__ count, bf16* __restrict__ dst, int C) {
  const int n = *count;
  for (int i = blockIdx.x; i < n; i += gridDim.x) {
    const uint4* s4 = reinterpret_cast<const uint4*>(src + (long long)rows[i] * C);
    uint4* d4 = reinterpret_cast<uint4*>(dst + (long long)i * C);
    for (int c = threadIdx.x; c < C / 8; c += blockDim.x) d4[c] = s4[c];
  }
}

// dst [M, C] = 0, then dst[rows[i], :] = src[i, :] for i < *count.  Block = one destination row; the inverse map is
// found by binary search in the (ascending) rows list.
__global__ void __launch_bounds__(256) scatter_rows_kernel(const bf16* __restrict__ src, const int* __restrict__ rows,
                                                           const int* __restrict__ count, bf16* __restrict__ dst, int M, int C) {
  const int n = *count;
  for (int m = blockIdx.x; m < M; m += gridDim.x) {
    int lo = 0, hi = n;
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (rows[mid] < m) lo = mid + 1; else hi = mid; }
    const bool hit = lo < n && rows[lo] == m;
    uint4* d4 = reinterpret_cast<uint4*>(dst + (long long)m * C);
    const uint4* s4 = reinterpret_cast<const uint4*>(src + (long long)lo * C);
    const uint4 z = make_uint4(0, 0, 0, 0);
    for (int c = threadIdx.x; c < C / 8; c += blockDim.x) d4[c] = hit ? s4[c] : z;
  }
}

}  // namespace dh

using namespace dh;

extern "C" int dh_rmsnorm_fwd(const void* x, int x_dtype, const float* weight, void* y, float* rstd, int M, int C, float eps,
                              void* stream) {
  DH_CHECK(x && weight && y && M > 0, "dh_rmsnorm_fwd: null argument");
  DH_CHECK(C % 256 == 0 && C <= 8192, "dh_rmsnorm_fwd: C=%d must be a multiple of 256 and <= 8192", C);
  cudaStream_t s = (cudaStream_t)stream;
  if (x_dtype == 0) DH_CUDA(launch_rmsnorm_fwd((const float*)x, weight, (bf16*)y, rstd, M, C, eps, s));
  else DH_CUDA(launch_rmsnorm_fwd((const bf16*)x, weight, (bf16*)y, rstd, M, C, eps, s));
  return 0;
}

extern "C" int dh_splitk_resid_rmsnorm(const float* partial, int splits, long long split_stride, const float* resid, const float* weight,
                                       float* x_out, void* y, int M, int C, float eps, void* stream) {
  DH_CHECK(partial && resid && weight && x_out && y && M > 0 && splits > 0, "dh_splitk_resid_rmsnorm: bad argument");
  DH_CHECK(C % 256 == 0 && C <= 8192, "dh_splitk_resid_rmsnorm: C=%d must be a multiple of 256 and <= 8192", C);
  DH_CHECK(split_stride >= (long long)M * C && split_stride % 4 == 0, "dh_splitk_resid_rmsnorm: bad split stride");
  DH_CUDA(launch_pdl(splitk_resid_rmsnorm_kernel, dim3(M), dim3(C / 8), 0, (cudaStream_t)stream, partial, splits, split_stride, resid, weight, x_out,
                     (bf16*)y, C, eps));
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_splitk_qkv_rope_cache(const float* partial, int splits, long long split_stride, void* qkv, void* cache_k, void* cache_v,
                                        const int32_t* seq, const int32_t* pos, const void* rope_cos, const void* rope_sin, int M, int max_len,
                                        int n_groups, int q_per_kv, void* stream) {
  DH_CHECK(partial && qkv && cache_k && cache_v && seq && pos && rope_cos && rope_sin && M > 0 && splits > 0,
           "dh_splitk_qkv_rope_cache: bad argument");
  const int N = n_groups * (q_per_kv + 2) * 64;
  DH_CHECK(split_stride >= (long long)M * N && split_stride % 4 == 0, "dh_splitk_qkv_rope_cache: bad split stride");
  int threads = N / 16;
  threads = threads > 256 ? 256 : ((threads + 31) / 32) * 32;
  DH_CUDA(launch_pdl(splitk_qkv_rope_cache_kernel, dim3(M), dim3(threads), 0, (cudaStream_t)stream, partial, splits, split_stride, (bf16*)qkv,
                     (bf16*)cache_k, (bf16*)cache_v, seq, pos, (const bf16*)rope_cos, (const bf16*)rope_sin, N, max_len, n_groups, q_per_kv));
  return 0;
}

extern "C" int dh_rmsnorm_bwd(const void* x, int x_dtype, const float* weight, const float* rstd, const void* dy,
                              const float* dres, float* dx_out, void* dx_bf16, int M, int C, void* stream) {
  DH_CHECK(x && weight && rstd && dy && dx_out && M > 0, "dh_rmsnorm_bwd: null argument");
  DH_CHECK(C % 256 == 0 && C <= 8192, "dh_rmsnorm_bwd: C=%d must be a multiple of 256 and <= 8192", C);
  cudaStream_t s = (cudaStream_t)stream;
  if (x_dtype == 0)
    DH_CUDA(launch_pdl(rmsnorm_bwd_kernel<float>, dim3(M), dim3(C / 8), 0, s, (const float*)x, weight, rstd, (const bf16*)dy, dres, dx_out, (bf16*)dx_bf16, C));
  else
    DH_CUDA(launch_pdl(rmsnorm_bwd_kernel<bf16>, dim3(M), dim3(C / 8), 0, s, (const bf16*)x, weight, rstd, (const bf16*)dy, dres, dx_out, (bf16*)dx_bf16, C));
  return 0;
}

extern "C" int dh_embedding_fwd(const int64_t* idx, const void* wte, int wte_dtype, void* out, int out_dtype, int M, int C,
                                int vocab_rows, void* stream) {
  DH_CHECK(idx && wte && out && M > 0 && C % 8 == 0, "dh_embedding_fwd: bad argument");
  cudaStream_t s = (cudaStream_t)stream;
  const int block = C / 8 < 256 ? ((C / 8 + 31) / 32) * 32 : 256;
  const long long* ix = (const long long*)idx;
  if (wte_dtype == 0 && out_dtype == 0) embedding_kernel<float, float><<<M, block, 0, s>>>(ix, (const float*)wte, (float*)out, C, vocab_rows);
  else if (wte_dtype == 0 && out_dtype == 1) embedding_kernel<float, bf16><<<M, block, 0, s>>>(ix, (const float*)wte, (bf16*)out, C, vocab_rows);
  else if (wte_dtype == 1 && out_dtype == 0) embedding_kernel<bf16, float><<<M, block, 0, s>>>(ix, (const bf16*)wte, (float*)out, C, vocab_rows);
  else embedding_kernel<bf16, bf16><<<M, block, 0, s>>>(ix, (const bf16*)wte, (bf16*)out, C, vocab_rows);
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_rope_apply(const void* x, void* out, const void* cos, const void* sin, int cs_dtype, int B, int T, int H, int D,
                             int rot, int interleaved, int conj, void* stream) {
  DH_CHECK(x && out && cos && sin, "dh_rope_apply: null argument");
  DH_CHECK(rot % 2 == 0 && rot <= D && rot > 0, "dh_rope_apply: rotary_dim %d must be even and <= headdim %d", rot, D);
  cudaStream_t s = (cudaStream_t)stream;
  const long long heads = (long long)B * T * H;
  const long long pairs = heads * (rot / 2);
  if (cs_dtype == 0)
    rope_apply_kernel<float><<<grid_for(pairs, 256), 256, 0, s>>>((const bf16*)x, (bf16*)out, (const float*)cos, (const float*)sin, pairs, T, H, D, rot, interleaved, conj);
  else
    rope_apply_kernel<bf16><<<grid_for(pairs, 256), 256, 0, s>>>((const bf16*)x, (bf16*)out, (const bf16*)cos, (const bf16*)sin, pairs, T, H, D, rot, interleaved, conj);
  if (out != x && rot < D) copy_tail_kernel<<<grid_for(heads * (D - rot), 256), 256, 0, s>>>((const bf16*)x, (bf16*)out, heads, D, rot);
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_rope_qkv_inplace(void* qkv, const void* cos, const void* sin, const int32_t* positions, int M, int seq_len,
                                   int n_groups, int q_per_kv, int conj, void* stream) {
  DH_CHECK(qkv && cos && sin && M > 0 && seq_len > 0, "dh_rope_qkv_inplace: bad argument");
  const long long total = (long long)M * n_groups * (q_per_kv + 2) * 4;
  rope_qkv_inplace_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>((bf16*)qkv, (const bf16*)cos, (const bf16*)sin,
                                                                                   positions, M, seq_len, n_groups, q_per_kv, conj);
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_cross_entropy(void* logits, int64_t ldl, const int64_t* targets, float* row_loss, int M, int V, float grad_scale,
                                int write_grad, void* stream) {
  DH_CHECK(logits && targets && row_loss && M > 0, "dh_cross_entropy: null argument");
  DH_CHECK(V % 8 == 0 && ldl % 8 == 0 && V * 2 <= 200 * 1024, "dh_cross_entropy: V=%d must be a multiple of 8 and fit shared memory", V);
  static bool attr = false;
  if (!attr) {
    DH_CUDA(cudaFuncSetAttribute(cross_entropy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr = true;
  }
  cross_entropy_kernel<<<M, 256, (size_t)V * 2, (cudaStream_t)stream>>>((bf16*)logits, ldl, (const long long*)targets, row_loss, V,
                                                                        grad_scale, write_grad);
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_loss_reduce(const float* row_loss, const int64_t* targets, float* out2, int M, void* stream) {
  DH_CHECK(row_loss && targets && out2 && M > 0, "dh_loss_reduce: null argument");
  loss_reduce_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(row_loss, (const long long*)targets, out2, M);
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_argmax(const void* logits, int64_t ldl, int64_t* ids, int M, int V, void* stream) {
  DH_CHECK(logits && ids && M > 0 && V > 0, "dh_argmax: bad argument");
  DH_CUDA(launch_pdl(argmax_kernel, dim3(M), dim3(256), 0, (cudaStream_t)stream, (const bf16*)logits, ldl, (long long*)ids, V));
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_cast_f32_to_bf16(const float* src, int64_t lds, void* dst, int64_t ldd, int64_t rows, int64_t cols, void* stream) {
  DH_CHECK(src && dst && rows > 0 && cols > 0 && cols % 8 == 0 && lds % 4 == 0 && ldd % 8 == 0, "dh_cast_f32_to_bf16: bad argument");
  cast_f32_bf16_kernel<<<grid_for(rows * cols / 8, 256), 256, 0, (cudaStream_t)stream>>>(src, lds, (bf16*)dst, ldd, rows, cols);
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_dropout_bf16(const void* x, void* out, int64_t n, float p, uint64_t seed, uint64_t offset, int accumulate, void* stream) {
  DH_CHECK(x && out && n > 0 && n % 8 == 0, "dh_dropout_bf16: n must be a positive multiple of 8");
  DH_CHECK(p >= 0.f && p < 1.f, "dh_dropout_bf16: p = %f not in [0, 1)", (double)p);
  DH_CHECK((((uintptr_t)x | (uintptr_t)out) & 15) == 0, "dh_dropout_bf16: operands must be 16-byte aligned");
  const uint32_t thresh = (uint32_t)lrintf(p * 65536.f);
  const float inv_keep = 65536.f / (float)(65536u - thresh);         // the keep probability the 16-bit threshold really gives
  const uint2 key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)), off = make_uint2((uint32_t)offset, (uint32_t)(offset >> 32));
  if (accumulate) dropout_bf16_kernel<true><<<grid_for(n / 8, 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)x, (bf16*)out, n / 8, thresh, inv_keep, key, off);
  else dropout_bf16_kernel<false><<<grid_for(n / 8, 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)x, (bf16*)out, n / 8, thresh, inv_keep, key, off);
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_adamw_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr, float beta1,
                             float beta2, float eps, float weight_decay, int step, float grad_scale, void* stream) {
  DH_CHECK(param && grad && exp_avg && exp_avg_sq && n > 0 && step > 0, "dh_adamw_step: bad argument");
  const float bc1 = 1.f - powf(beta1, (float)step);
  const float bc2s = sqrtf(1.f - powf(beta2, (float)step));
  adamw_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(param, grad, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps,
                                                                    weight_decay, bc1, bc2s, grad_scale);
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_kv_cache_write(const void* qkv, void* cache_k, void* cache_v, const int32_t* seq, const int32_t* pos, int M,
                                 int max_len, int n_groups, int q_per_kv, void* stream) {
  DH_CHECK(qkv && cache_k && cache_v && seq && pos && M > 0, "dh_kv_cache_write: null argument");
  const long long total = (long long)M * n_groups * 16;
  kv_cache_write_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)qkv, (bf16*)cache_k, (bf16*)cache_v, seq,
                                                                                 pos, M, max_len, n_groups, q_per_kv);
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_prefix_project_fwd(const void* prefix, const void* wkv, void* prefix_kv, int L, int aT, int C, int n_groups,
                                     int q_per_kv, void* stream) {
  DH_CHECK(prefix && wkv && prefix_kv && L > 0 && aT > 0 && aT <= PFX_MAX && C % 256 == 0, "dh_prefix_project_fwd: bad argument");
  const int R = n_groups * 128;
  const size_t smem = (size_t)aT * C * 2;
  DH_CHECK(smem <= 96 * 1024, "dh_prefix_project_fwd: prefix rows do not fit in shared memory");
  static bool attr = false;
  if (!attr) {
    DH_CUDA(cudaFuncSetAttribute(prefix_project_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    attr = true;
  }
  const int rows_per_block = 32;
  const long long ld_out = (long long)n_groups * (q_per_kv + 2) * 64;
  prefix_project_fwd_kernel<<<dim3((R + rows_per_block - 1) / rows_per_block, L), 256, smem, (cudaStream_t)stream>>>(
      (const bf16*)prefix, (const bf16*)wkv, (bf16*)prefix_kv, aT, C, R, n_groups, q_per_kv, ld_out, rows_per_block);
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_prefix_project_bwd(const float* d_prefix_kv, const void* wkv, float* d_prefix, int L, int aT, int C, int n_groups,
                                     int q_per_kv, void* stream) {
  DH_CHECK(d_prefix_kv && wkv && d_prefix && L > 0 && aT > 0 && aT <= PFX_MAX && C % 256 == 0, "dh_prefix_project_bwd: bad argument");
  const int R = n_groups * 128;
  const size_t smem = (size_t)aT * R * 4;
  DH_CHECK(smem <= 48 * 1024, "dh_prefix_project_bwd: too many kv rows");
  const long long ld = (long long)n_groups * (q_per_kv + 2) * 64;
  prefix_project_bwd_kernel<<<dim3(C / 256, L), 128, smem, (cudaStream_t)stream>>>(d_prefix_kv, (const bf16*)wkv, d_prefix, aT, C, R,
                                                                                   n_groups, q_per_kv, ld);
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_compact_targets(const int64_t* targets, int M, int32_t* count, int32_t* rows, int64_t* targets_c, void* stream) {
  DH_CHECK(targets && count && rows && targets_c && M > 0, "dh_compact_targets: bad argument");
  compact_targets_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>((const long long*)targets, M, count, rows, (long long*)targets_c);
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_gather_rows(const void* src, const int32_t* rows, const int32_t* count, void* dst, int M, int C, void* stream) {
  DH_CHECK(src && rows && count && dst && M > 0 && C > 0 && C % 8 == 0, "dh_gather_rows: bad argument");
  gather_rows_kernel<<<M < 1184 ? M : 1184, 256, 0, (cudaStream_t)stream>>>((const bf16*)src, rows, count, (bf16*)dst, C);
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_scatter_rows(const void* src, const int32_t* rows, const int32_t* count, void* dst, int M, int C, void* stream) {
  DH_CHECK(src && rows && count && dst && M > 0 && C > 0 && C % 8 == 0, "dh_scatter_rows: bad argument");
  scatter_rows_kernel<<<M < 2368 ? M : 2368, 256, 0, (cudaStream_t)stream>>>((const bf16*)src, rows, count, (bf16*)dst, M, C);
  DH_CUDA(cudaGetLastError());
  return 0;
}
