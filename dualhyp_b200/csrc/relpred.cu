// RelPrompt reliability predictor head (ger/relprompt.py:126-147) forward and backward, fp32:
//   Conv1d(C->256,k3,p1)+ReLU -> Conv1d(256->256,k3,p1)+ReLU -> AvgPool1d(pool, ceil_mode) -> Linear(256->3)
// The three convolution-shaped contractions (forward, input gradient, weight gradient) share one tiled
// implicit-GEMM kernel whose operand fetch and output store are supplied by small problem structs, so no
// im2col buffer is ever materialised.  Sizes are tiny next to the decoder (about 4 GFLOP per step); this
// stays on the fp32 CUDA-core pipe because the reference runs the head in fp32.
#include "common.cuh"
#include "../../include/dualhyp.h"

namespace dh {

constexpr int HID = 256;

// C[i,j] = sum_k A(i,k) * B(k,j);   64x64 tile, 16-deep k slices, 256 threads, 4x4 outputs per thread.
template <class P>
__global__ void __launch_bounds__(256) igemm_kernel(const P prob) {
  __shared__ float As[16][64 + 4];
  __shared__ float Bs[16][64 + 4];
  const int i0 = blockIdx.y * 64, j0 = blockIdx.x * 64;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
  for (int k0 = 0; k0 < prob.K; k0 += 16) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int lin = threadIdx.x + e * 256;   // 1024 elements per operand slice
      const int kk = lin & 15, ij = lin >> 4;
      const int k = k0 + kk;
      As[kk][ij] = (k < prob.K && i0 + ij < prob.I) ? prob.a(i0 + ij, k) : 0.f;
      const int kk2 = lin >> 6, jj = lin & 63;
      const int k2 = k0 + kk2;
      Bs[kk2][jj] = (k2 < prob.K && j0 + jj < prob.J) ? prob.b(k2, j0 + jj) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float av[4], bv[4];
#pragma unroll
      for (int a = 0; a < 4; ++a) av[a] = As[kk][ty * 4 + a];
#pragma unroll
      for (int b = 0; b < 4; ++b) bv[b] = Bs[kk][tx * 4 + b];
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] += av[a] * bv[b];
    }
    __syncthreads();
  }
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const int i = i0 + ty * 4 + a, j = j0 + tx * 4 + b;
      if (i < prob.I && j < prob.J) prob.store(i, j, acc[a][b]);
    }
}

// forward conv: i = (b,t), j = out channel, k = tap*Cin + c.   out = relu(bias + conv)
struct ConvFwd {
  int I, J, K, T, Cin;
  const float *x, *w, *bias;
  float* out;
  __device__ float a(int i, int k) const {
    const int tap = k / Cin, c = k - tap * Cin;
    const int t = i % T + tap - 1;
    return (t >= 0 && t < T) ? x[(long long)(i - i % T + t) * Cin + c] : 0.f;
  }
  __device__ float b(int k, int j) const {
    const int tap = k / Cin, c = k - tap * Cin;
    return w[((long long)j * Cin + c) * 3 + tap];
  }
  __device__ void store(int i, int j, float v) const { out[(long long)i * J + j] = fmaxf(v + bias[j], 0.f); }
};

// input gradient of conv2 through ReLU of h1: i = (b,t), j = in channel c, k = tap*HID + o
//   dz1[b,t,c] = (h1 > 0) * sum_{tap,o} dz2[b, t - tap + 1, o] * w2[o, c, tap]
struct ConvDgrad {
  int I, J, K, T;
  const float *dz, *w, *h_in;
  float* out;
  __device__ float a(int i, int k) const {
    const int tap = k / HID, o = k - tap * HID;
    const int t = i % T - tap + 1;
    return (t >= 0 && t < T) ? dz[(long long)(i - i % T + t) * HID + o] : 0.f;
  }
  __device__ float b(int k, int j) const {
    const int tap = k / HID, o = k - tap * HID;
    return w[((long long)o * J + j) * 3 + tap];
  }
  __device__ void store(int i, int j, float v) const { out[(long long)i * J + j] = h_in[(long long)i * J + j] > 0.f ? v : 0.f; }
};

// weight gradient: i = out channel o, j = c*3 + tap (PyTorch weight layout), k = (b,t)
//   dw[o,c,tap] += sum_{b,t} dz[b,t,o] * in[b, t + tap - 1, c]
struct ConvWgrad {
  int I, J, K, T, Cin;
  const float *dz, *in;
  float* dw;
  __device__ float a(int i, int k) const { return dz[(long long)k * HID + i]; }
  __device__ float b(int k, int j) const {
    const int c = j / 3, tap = j - c * 3;
    const int t = k % T + tap - 1;
    return (t >= 0 && t < T) ? in[(long long)(k - k % T + t) * Cin + c] : 0.f;
  }
  __device__ void store(int i, int j, float v) const { dw[(long long)i * J + j] += v; }
};

// AvgPool1d(kernel=stride=pool, ceil_mode=True): the last window averages its valid frames only.
__global__ void pool_fwd_kernel(const float* __restrict__ h2, float* __restrict__ pooled, int B, int T, int pool, int n_out) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * n_out * HID) return;
  const int c = idx % HID, jo = (idx / HID) % n_out, b = idx / (HID * n_out);
  const int t0 = jo * pool, t1 = min(T, t0 + pool);
  float s = 0.f;
  for (int t = t0; t < t1; ++t) s += h2[((long long)b * T + t) * HID + c];
  pooled[idx] = s / (float)(t1 - t0);
}
__global__ void classifier_fwd_kernel(const float* __restrict__ pooled, const float* __restrict__ wc, const float* __restrict__ bc,
                                      float* __restrict__ logits, int rows) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= rows * 3) return;
  const int r = warp / 3, k = warp % 3;
  float s = 0.f;
  for (int c = lane; c < HID; c += 32) s += pooled[(long long)r * HID + c] * wc[k * HID + c];
  s = warp_sum(s);
  if (lane == 0) logits[r * 3 + k] = s + bc[k];
}
// dz2[b,t,c] = (h2 > 0) * (sum_k dlogits[b,jo,k] * wc[k,c]) / window_len ;  also dwc, dbc (one block per class row set)
__global__ void pool_bwd_kernel(const float* __restrict__ h2, const float* __restrict__ dlogits, const float* __restrict__ wc,
                                float* __restrict__ dz2, int B, int T, int pool, int n_out) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)B * T * HID) return;
  const int c = (int)(idx % HID);
  const int t = (int)((idx / HID) % T), b = (int)(idx / ((long long)HID * T));
  const int jo = t / pool;
  const int wl = min(T, (jo + 1) * pool) - jo * pool;
  const float* dl = dlogits + ((long long)b * n_out + jo) * 3;
  const float dp = dl[0] * wc[c] + dl[1] * wc[HID + c] + dl[2] * wc[2 * HID + c];
  dz2[idx] = h2[idx] > 0.f ? dp / (float)wl : 0.f;
}
__global__ void classifier_bwd_kernel(const float* __restrict__ pooled, const float* __restrict__ dlogits, float* dwc, float* dbc, int rows) {
  // one thread per (k, c); serial over rows (rows is a few dozen)
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < 3 * HID) {
    const int k = idx / HID, c = idx % HID;
    float s = 0.f;
    for (int r = 0; r < rows; ++r) s += dlogits[r * 3 + k] * pooled[(long long)r * HID + c];
    dwc[idx] += s;
  } else if (idx < 3 * HID + 3) {
    const int k = idx - 3 * HID;
    float s = 0.f;
    for (int r = 0; r < rows; ++r) s += dlogits[r * 3 + k];
    dbc[k] += s;
  }
}
// db[o] += sum_rows dz[row, o]
__global__ void bias_grad_kernel(const float* __restrict__ dz, float* db, int rows) {
  const int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= HID) return;
  float s = 0.f;
  for (int r = 0; r < rows; ++r) s += dz[(long long)r * HID + o];
  db[o] += s;
}

template <class P>
static int run_igemm(const P& prob, cudaStream_t s) {
  dim3 grid((prob.J + 63) / 64, (prob.I + 63) / 64);
  igemm_kernel<P><<<grid, 256, 0, s>>>(prob);
  DH_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace dh

using namespace dh;

extern "C" int dh_relpred_fwd(const float* x, const float* w1, const float* b1, const float* w2, const float* b2, const float* wc,
                              const float* bc, float* h1, float* h2, float* pooled, float* logits, int B, int T, int C, int pool,
                              void* stream) {
  DH_CHECK(x && w1 && b1 && w2 && b2 && wc && bc && h1 && h2 && pooled && logits, "dh_relpred_fwd: null argument");
  DH_CHECK(B > 0 && T > 0 && C > 0 && pool > 0, "dh_relpred_fwd: bad shape");
  cudaStream_t s = (cudaStream_t)stream;
  const int n_out = (T + pool - 1) / pool;
  ConvFwd c1 = {B * T, HID, 3 * C, T, C, x, w1, b1, h1};
  if (run_igemm(c1, s)) return -2;
  ConvFwd c2 = {B * T, HID, 3 * HID, T, HID, h1, w2, b2, h2};
  if (run_igemm(c2, s)) return -2;
  pool_fwd_kernel<<<(B * n_out * HID + 255) / 256, 256, 0, s>>>(h2, pooled, B, T, pool, n_out);
  classifier_fwd_kernel<<<(B * n_out * 3 * 32 + 255) / 256, 256, 0, s>>>(pooled, wc, bc, logits, B * n_out);
  DH_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dh_relpred_bwd(const float* x, const float* w1, const float* w2, const float* wc, const float* h1, const float* h2,
                              const float* pooled, const float* d_logits, float* dw1, float* db1, float* dw2, float* db2, float* dwc,
                              float* dbc, float* scratch, int B, int T, int C, int pool, void* stream) {
  DH_CHECK(x && w1 && w2 && wc && h1 && h2 && pooled && d_logits && dw1 && db1 && dw2 && db2 && dwc && dbc && scratch,
           "dh_relpred_bwd: null argument");
  cudaStream_t s = (cudaStream_t)stream;
  const int n_out = (T + pool - 1) / pool;
  const int rows = B * T;
  float* dz2 = scratch;                          // [B*T, 256]
  float* dz1 = scratch + (long long)rows * HID;  // [B*T, 256]
  classifier_bwd_kernel<<<(3 * HID + 3 + 255) / 256, 256, 0, s>>>(pooled, d_logits, dwc, dbc, B * n_out);
  pool_bwd_kernel<<<(int)(((long long)rows * HID + 255) / 256), 256, 0, s>>>(h2, d_logits, wc, dz2, B, T, pool, n_out);
  bias_grad_kernel<<<1, HID, 0, s>>>(dz2, db2, rows);
  ConvWgrad g2 = {HID, HID * 3, rows, T, HID, dz2, h1, dw2};
  if (run_igemm(g2, s)) return -2;
  ConvDgrad d1 = {rows, HID, 3 * HID, T, dz2, w2, h1, dz1};
  if (run_igemm(d1, s)) return -2;
  bias_grad_kernel<<<1, HID, 0, s>>>(dz1, db1, rows);
  ConvWgrad g1 = {HID, C * 3, rows, T, C, dz1, x, dw1};
  if (run_igemm(g1, s)) return -2;
  DH_CUDA(cudaGetLastError());
  return 0;
}

// The pooling + classifier tail of the head on its own (the tensor-core path runs the two convolutions as bf16 GEMMs through
// dh_gemm -- rows of three consecutive frames are the K-major A operand, no im2col kernel -- and only needs the tail here).
extern "C" int dh_relpred_head_fwd(const float* h2, const float* wc, const float* bc, float* pooled, float* logits, int B, int T, int pool,
                                   void* stream) {
  DH_CHECK(h2 && wc && bc && pooled && logits && B > 0 && T > 0 && pool > 0, "dh_relpred_head_fwd: bad argument");
  cudaStream_t s = (cudaStream_t)stream;
  const int n_out = (T + pool - 1) / pool;
  pool_fwd_kernel<<<(B * n_out * HID + 255) / 256, 256, 0, s>>>(h2, pooled, B, T, pool, n_out);
  classifier_fwd_kernel<<<(B * n_out * 3 * 32 + 255) / 256, 256, 0, s>>>(pooled, wc, bc, logits, B * n_out);
  DH_CUDA(cudaGetLastError());
  return 0;
}

// d_logits -> dz2 = d(conv2 pre-activation) f32 [B, T, 256] (through the average pool and the ReLU of h2); dwc / dbc ACCUMULATED.
extern "C" int dh_relpred_head_bwd(const float* h2, const float* pooled, const float* d_logits, const float* wc, float* dz2, float* dwc,
                                   float* dbc, int B, int T, int pool, void* stream) {
  DH_CHECK(h2 && pooled && d_logits && wc && dz2 && dwc && dbc && B > 0 && T > 0 && pool > 0, "dh_relpred_head_bwd: bad argument");
  cudaStream_t s = (cudaStream_t)stream;
  const int n_out = (T + pool - 1) / pool;
  classifier_bwd_kernel<<<(3 * HID + 3 + 255) / 256, 256, 0, s>>>(pooled, d_logits, dwc, dbc, B * n_out);
  pool_bwd_kernel<<<(int)(((long long)B * T * HID + 255) / 256), 256, 0, s>>>(h2, d_logits, wc, dz2, B, T, pool, n_out);
  DH_CUDA(cudaGetLastError());
  return 0;
}
