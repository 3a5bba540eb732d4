/*
 * dualhyp.h -- C ABI of libdualhyp.so: the B200 (sm_100a) device path of the DualHyp decoder hot path.
 *
 * The reference (sungnyun/dualhyp) is pure Python over PyTorch; its hot path has no FFI of its own
 * (SURVEY.md section 8b).  These entry points are what the reference-side binding -- a ctypes stub inside
 * the `ger` package, shown in INTEGRATION.md and shipped as dualhyp_b200/_lib.py -- binds in place of the
 * ATen / cuBLAS / flash-attn calls the reference makes.  Each entry cites the reference site it replaces.
 *
 * Conventions: plain pointers and sizes only; every pointer is a DEVICE pointer unless stated; the library
 * never allocates or frees caller-visible memory and keeps no mutable global state except a per-device
 * TMA-descriptor cache, the last-error string and the static scratch array of dh_gemm_swiglu_streamk_noalloc; all work is enqueued on `stream` (a cudaStream_t) and is
 * stream-ordered; functions return 0 on success, negative on error (dh_last_error() tells why).
 * bf16 = 2-byte bfloat16.  Row-major everywhere; `ld*` are leading dimensions in ELEMENTS.
 */
#ifndef DUALHYP_H
#define DUALHYP_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DH_ABI_VERSION 4

int dh_abi_version(void);
const char* dh_last_error(void);
/* Idempotent per-device initialisation (function attributes, SM count). Fails when no sm_100 device. */
int dh_init(int device);
int dh_sm_count(void);

/* ---------------------------------------------------------------------------------------------------
 * GEMM core (tcgen05.mma + TMEM accumulators + TMA operand staging), replaces every nn.Linear on the path:
 * ger/model.py:213,266,313-316 ; ger/lora.py:162-166,388-402 ; ger/lora.py:70-71 (lm_head).
 *   D[M,N] = A[M,K1] * op(B) (+ A2[M,K2] * op(B2)),   bf16 operands, fp32 accumulation in TMEM.
 * a_layout/b_layout 0 = K-major  (A is [M,K] row-major; B is [N,K] row-major, i.e. y = x W^T with W as stored)
 *                   1 = MN-major (A is [K,M] row-major; B is [K,N] row-major)            -- dgrad / wgrad forms
 * b_layout may carry DH_B_FROZEN: B is a frozen weight that no kernel on the stream writes.  The decode GEMM (M <= 128) then
 * requests its first ring of weight tiles before waiting for the previous kernel (programmatic dependent launch, DH_PDL=2),
 * so the weight stream of one GEMM overlaps the tail of the kernel before it.  Never set it for an operand produced on the stream.
 * The second source pair folds a LoRA low-rank delta into the same accumulator (K extension).
 * Constraints: N % 64 == 0, 16-byte aligned rows (so K % 8 == 0 for a K-major operand and M % 8 / N % 8 for an MN-major one).
 * M is arbitrary and K may be ragged (tails are zero-filled by TMA on load; tail rows are masked on store); with both operands
 * MN-major (the wgrad form, K = token count of a packed batch) K is arbitrary.
 * ------------------------------------------------------------------------------------------------- */
#define DH_B_FROZEN 0x100

enum dh_epilogue {
  DH_EPI_BF16 = 0,        /* out[bf16] = (acc*alpha [+col_bias]) [*col_scale]                                  */
  DH_EPI_F32 = 1,         /* out[f32]  = acc*alpha (+ out if accumulate)                                       */
  DH_EPI_RESID_F32 = 2,   /* out[f32]  = aux[f32] + bf16(acc)        residual add, fp32 stream (ger/model.py:185-186) */
  DH_EPI_RESID_BF16 = 3,  /* out[bf16] = bf16(aux[bf16] + bf16(acc)) residual add, bf16 stream (bf16-true inference)  */
  DH_EPI_QKV_ROPE = 4,    /* out[bf16] = NeoX RoPE on the q and k head slots of the group-interleaved QKV row,
                             v copied (ger/model.py:216-246,349-355); positions[row] or row % seq_len           */
  DH_EPI_SWIGLU = 5,      /* columns are [fc_1 x32 | fc_2 x32] interleaved blocks: out[bf16, ld N] = raw h12 (may be
                             NULL), out2[bf16, ld N/2] = silu(h1)*h2  (ger/model.py:313-316)                   */
  DH_EPI_SWIGLU_BWD = 6,  /* acc = d(act) [M, N]; aux = h12 [M, 2N] interleaved; out[bf16, 2N] = d(h12)          */
  DH_EPI_ROPE_BWD_QKV = 7 /* reserved */
};

typedef struct dh_gemm_args {
  const void* A;  int64_t lda;
  const void* B;  int64_t ldb;
  const void* A2; int64_t lda2;      /* optional (K2 == 0 -> unused) */
  const void* B2; int64_t ldb2;
  int32_t M, N, K1, K2;
  int32_t a_layout, b_layout, a2_layout, b2_layout;
  int32_t epilogue;
  int32_t accumulate;                /* DH_EPI_F32 only */
  float alpha;
  int32_t seq_len;                   /* DH_EPI_QKV_ROPE: row -> position = row % seq_len when positions == NULL */
  int32_t q_per_kv;                  /* DH_EPI_QKV_ROPE: q heads per kv head (group = q_per_kv + 2 slots of 64)   */
  int32_t split_k;                   /* > 1 (DH_EPI_F32 only, no accumulate, no m_limit): `out` is fp32 [split_k][M][ldo]; K range s of
                                        split_k is multiplied into slice s and NO reduction is done -- the consumer sums the slices
                                        (dh_splitk_resid_rmsnorm in the decode step; a plain sum for the LoRA wgrads whose K is the
                                        token count).  0 / 1: off. */
  void* out;  int64_t ldo;
  void* out2; int64_t ldo2;
  const void* aux; int64_t ldaux;
  const float* col_scale;            /* optional [N]  (AdapterV2Linear.adapter_scale, ger/lora.py:70-71) */
  const float* col_bias;             /* optional [N]  (AdapterV2Linear.adapter_bias) */
  const void* rope_cos;              /* bf16 [block_size, 64] as built by ger/model.py:319-346 */
  const void* rope_sin;
  const int32_t* positions;          /* optional [M] */
  const int32_t* m_limit;            /* optional device int: only rows [0, *m_limit) are needed; whole tiles beyond are skipped
                                        (rows of a partly needed tile are still computed).  ABI v2. */
  void* workspace;                   /* optional, ABI v4: device memory the call may use, ZEROED once by the owner and only ever passed to
                                        calls of the same (M-class, N, K).  With it a decode GEMM (M <= 256) whose epilogue is tile-local
                                        (DH_EPI_SWIGLU) and whose N / 128 tiles do not fill the SMs runs stream-K over ALL SMs: partial
                                        tiles + per-tile arrival counters live here, the CTA completing a tile runs the epilogue. */
  size_t workspace_bytes;            /* >= dh_gemm_streamk_workspace_bytes(M, N, K1) */
} dh_gemm_args;

int dh_gemm(const dh_gemm_args* args, void* stream);
size_t dh_gemm_streamk_workspace_bytes(int M, int N, int K);

/* No counterpart in the reference: added for callers that may not allocate (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The stream-K DH_EPI_SWIGLU decode GEMM of dh_gemm
 * (args->workspace given) with the partial tiles and arrival counters in a static device array of DH_NOALLOC_SCRATCH_BYTES that
 * every GPU loading the library holds until exit; args->workspace / workspace_bytes are ignored.  The call makes no cudaMalloc,
 * cudaFree or allocating library call.  Accepted: epilogue DH_EPI_SWIGLU, M <= 256, N % 128 == 0, K1 % 64 == 0, K2 == 0, K-major
 * operands, no m_limit, N / 128 < SM count, and dh_gemm_streamk_workspace_bytes(M, N, K1) <= DH_NOALLOC_SCRATCH_BYTES (TinyLlama's
 * fc_1|fc_2 at 256 rows needs 33.0 MiB); anything else returns an error that dh_last_error() explains -- it never falls back.
 * The arrival counters are cleared on the stream before every launch.  ALL calls on a device share the one array: calls that
 * may overlap on the GPU (two streams, two threads) race; serialise them.  The first call on a device looks the array's address
 * up (cudaGetSymbolAddress); make it outside stream capture. */
#define DH_NOALLOC_SCRATCH_BYTES ((size_t)36 << 20)
int dh_gemm_swiglu_streamk_noalloc(const dh_gemm_args* args, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * Bandwidth-bound leaf ops
 * ------------------------------------------------------------------------------------------------- */
/* ger/rmsnorm.py:17-21.  x [M,C] (x_dtype 0 = f32, 1 = bf16), weight f32 [C]; y bf16 [M,C]; rstd f32 [M]
 * (may be NULL).  C % 8 == 0, C <= 8192. */
int dh_rmsnorm_fwd(const void* x, int x_dtype, const float* weight, void* y, float* rstd, int M, int C, float eps,
                   void* stream);
/* Decode: consumer of a split-K dh_gemm (split_k > 1).  partial f32 [splits][M][C] (slice pitch split_stride elements):
 *   x_out f32 [M,C] = resid + bf16(sum_s partial[s])   -- the residual add of ger/model.py:185-186, slices summed in order;
 *   y bf16 [M,C]    = rmsnorm(x_out) * weight          -- the norm that follows it (ger/rmsnorm.py:17-21).
 * x_out may alias resid. */
int dh_splitk_resid_rmsnorm(const float* partial, int splits, long long split_stride, const float* resid, const float* weight,
                            float* x_out, void* y, int M, int C, float eps, void* stream);
/* Decode: consumer of the split-K QKV dh_gemm.  partial f32 [splits][M][N], N = n_groups*(qpk+2)*64:
 *   qkv bf16 [M,N] = RoPE(bf16(sum_s partial[s])) on the q and k slots at pos[row] (the DH_EPI_QKV_ROPE math, ger/model.py:240-246),
 *   and the k / v slots are also written to the KV cache at (seq[row], pos[row]) (ger/model.py:248-259) -- one launch. */
int dh_splitk_qkv_rope_cache(const float* partial, int splits, long long split_stride, void* qkv, void* cache_k, void* cache_v,
                             const int32_t* seq, const int32_t* pos, const void* rope_cos, const void* rope_sin, int M, int max_len,
                             int n_groups, int q_per_kv, void* stream);
/* Backward of the above w.r.t. x (weights are frozen on this path).  dy bf16 [M,C].
 * dx_out f32 [M,C] = (dres ? dres : 0) + d(rmsnorm)/dx ; dx_bf16 (optional) = bf16(dx_out).
 * dres may alias dx_out. */
int dh_rmsnorm_bwd(const void* x, int x_dtype, const float* weight, const float* rstd, const void* dy,
                   const float* dres, float* dx_out, void* dx_bf16, int M, int C, void* stream);
/* nn.Embedding gather (ger/lora.py:537): out (dtype 0 = f32, 1 = bf16) [M,C] = wte[idx[m]] ; wte dtype likewise. */
int dh_embedding_fwd(const int64_t* idx, const void* wte, int wte_dtype, void* out, int out_dtype, int M, int C,
                     int vocab_rows, void* stream);
/* ger/fused_rotary_embedding.py:10-90 (apply_rotary_emb_func): x bf16 (B, T, H, D) contiguous, cos/sin f32 or bf16
 * [T, rot/2]; rotates the first `rot` dims of every head, NeoX halves (interleaved=0) or GPT-J pairs (1);
 * conj=1 is the backward rotation.  out may alias x (inplace). */
int dh_rope_apply(const void* x, void* out, const void* cos, const void* sin, int cs_dtype, int B, int T, int H, int D,
                  int rot, int interleaved, int conj, void* stream);
/* In-place NeoX RoPE (conj = backward) on the q and k slots of a group-interleaved [M, n_groups*(qpk+2)*64] bf16
 * buffer -- the backward of DH_EPI_QKV_ROPE. */
int dh_rope_qkv_inplace(void* qkv, const void* cos, const void* sin, const int32_t* positions, int M, int seq_len,
                        int n_groups, int q_per_kv, int conj, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * Attention (ger/model.py:270-290 causal SDPA with native GQA; ger/adapter.py:116-146 gated prefix)
 * qkv: bf16 [B*T, n_groups*(qpk+2)*64] group-interleaved, RoPE already applied.  out: bf16 [B*T, n_head*64].
 * lse: f32 [B, n_head, T] (log-sum-exp of the scaled scores, natural log).  head_size is 64.
 * Prefix (optional, prefix_kv != NULL): prefix_kv bf16 [aT, n_groups*(qpk+2)*64] (the qkv linear applied to
 * adapter_wte; only its k and v slots are read, no RoPE), gate f32 [n_head]:
 *     out = softmax_causal(q k^T / 8) v + gate[h] * softmax(q ak^T / 8) av
 * ------------------------------------------------------------------------------------------------- */
int dh_attention_fwd(const void* qkv, void* out, float* lse, const void* prefix_kv, const float* gate, int B, int T,
                     int n_groups, int q_per_kv, int prefix_len, void* stream);
/* The same forward over PACKED variable-length sequences (no pad rows; SURVEY 8f row 2: the prompts of a decode batch are
 * 330-480 tokens of 480): sequence b = rows [cu_seqlens[b], cu_seqlens[b+1]) of qkv / out (total_rows rows), causal inside
 * each sequence; max_len = the longest sequence (sizes the grid); lse: f32 [n_head, total_rows].  RoPE positions restart
 * per sequence (dh_gemm `positions`). */
int dh_attention_fwd_varlen(const void* qkv, void* out, float* lse, const void* prefix_kv, const float* gate,
                            const int32_t* cu_seqlens, int n_seq, int max_len, long long total_rows, int n_groups, int q_per_kv,
                            int prefix_len, void* stream);
/* Backward.  dout bf16 [B*T, n_head*64].  dqkv bf16 (same layout as qkv; every slot is written).
 * d_prefix_kv f32 [aT, n_groups*(qpk+2)*64] and d_gate f32 [n_head] are ACCUMULATED into (caller zeroes).
 * need_dqkv = 0 computes only the prefix/gate gradients (first adapter layer). workspace: dh_attention_bwd_workspace.
 * rope_cos/rope_sin (bf16 [block_size, 64], or both NULL): when given, the conjugate RoPE rotation -- the backward of
 * DH_EPI_QKV_ROPE -- is fused into the dq and dk stores, so dqkv is the gradient w.r.t. the PRE-rotation QKV row
 * (position of row m is m % T). */
size_t dh_attention_bwd_workspace(int B, int T, int n_groups, int q_per_kv);
int dh_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, const void* prefix_kv,
                     const float* gate, void* dqkv, float* d_prefix_kv, float* d_gate, int need_dqkv,
                     const void* rope_cos, const void* rope_sin, int B, int T, int n_groups, int q_per_kv, int prefix_len,
                     void* workspace, size_t workspace_bytes, void* stream);
/* The same backward over PACKED variable-length sequences (training on DualHyp prompts of 330-480 tokens without the right
 * pad of data/av_dataset.py:258-292): sequence b = rows [cu_seqlens[b], cu_seqlens[b+1]) of qkv, out, dout and
 * dqkv (total_rows rows), causal inside each sequence; lse as written by dh_attention_fwd_varlen (f32 [n_head, total_rows]);
 * RoPE positions restart per sequence; workspace >= total_rows * n_head * 4 bytes.  ABI v3. */
int dh_attention_bwd_varlen(const void* qkv, const void* out, const void* dout, const float* lse, const void* prefix_kv,
                            const float* gate, void* dqkv, float* d_prefix_kv, float* d_gate, int need_dqkv,
                            const void* rope_cos, const void* rope_sin, const int32_t* cu_seqlens, int n_seq, int max_len,
                            long long total_rows, int n_groups, int q_per_kv, int prefix_len, void* workspace,
                            size_t workspace_bytes, void* stream);
/* Adapter prefix K/V for ALL adapter layers in one launch (ger/adapter.py:129-137, `self.attn(prefix)` of which only the k
 * and v slots are used).  prefix bf16 [L, aT, C]; wkv bf16 [L, n_groups*128, C] = the k rows (64) then the v rows (64) of
 * every group of each layer's QKV weight; prefix_kv bf16 [L, aT, n_groups*(qpk+2)*64]: only the k and v slots are written.
 * The backward takes d(prefix_kv) f32 in that same layout (what dh_attention_bwd accumulates) -> d_prefix f32 [L, aT, C]. */
int dh_prefix_project_fwd(const void* prefix, const void* wkv, void* prefix_kv, int L, int aT, int C, int n_groups, int q_per_kv,
                          void* stream);
int dh_prefix_project_bwd(const float* d_prefix_kv, const void* wkv, float* d_prefix, int L, int aT, int C, int n_groups,
                          int q_per_kv, void* stream);
/* Single-query attention over a KV cache (ger/model.py:248-261 decode step).  q bf16 [nseq, n_head*64] taken from
 * the q slots of qkv rows; cache_k/v bf16 [nseq, n_groups, max_len, 64]; lens int32 [nseq] = number of valid keys
 * INCLUDING the token being decoded.  out bf16 [nseq, n_head*64]. */
int dh_attention_decode(const void* qkv, const void* cache_k, const void* cache_v, const int32_t* lens, void* out,
                        const void* prefix_kv, const float* gate, int nseq, int max_len, int n_groups, int q_per_kv,
                        int prefix_len, void* stream);
/* The kernel dh_attention_decode uses by default when max_len is a multiple of 32 (csrc/decode_attn.cu): one persistent block per
 * SM walks its share of the (sequence, kv group) items, their cached K / V rows stream through a TMA ring as 32-key chunks,
 * flash-style running softmax per warp, same arguments; prefix_len <= 32.  lens, prefix_kv and gate are read before
 * griddepcontrol.wait: they must not be the output of a kernel that signals programmatic launch completion early (none of this
 * library's producers of them does; a kernel launched without that attribute never does).  Cache slots at or beyond lens[i]
 * may hold any FINITE values (their softmax weight is an exact zero, and 0 * NaN is not): allocate the cache zeroed.
 * dh_attention_decode_rows is the other one, explicitly: whole-row scores in shared memory on mma.sync tiles (what
 * DH_DEC_ATTN=mma selects, and the fallback for other row lengths). */
int dh_attention_decode_stream(const void* qkv, const void* cache_k, const void* cache_v, const int32_t* lens, void* out,
                               const void* prefix_kv, const float* gate, int nseq, int max_len, int n_groups, int q_per_kv,
                               int prefix_len, void* stream);
int dh_attention_decode_rows(const void* qkv, const void* cache_k, const void* cache_v, const int32_t* lens, void* out,
                             const void* prefix_kv, const float* gate, int nseq, int max_len, int n_groups, int q_per_kv,
                             int prefix_len, void* stream);
/* ---------------------------------------------------------------------------------------------------
 * One decode token step for ALL decoder layers as one persistent kernel (ger/model.py:163-268 per layer, driven one token at
 * a time by generate/base.py:57-80): one CTA per SM, grid-wide barriers between the stages, the frozen weights of every GEMM
 * streamed by a producer warp that runs ahead of the barriers (stream-K units over all SMs), split-K partial tiles summed by the
 * stage that follows (RoPE + KV-cache write inside the attention stage, residual + RMSNorm, SwiGLU).  n_seq <= 128 sequences.
 *   in : x f32 [n_seq, C] residual stream (the token embeddings), n bf16 [n_seq, C] = rmsnorm(x) * layer 0's norm_1 weight,
 *        positions int32 [n_seq] (position of the token being decoded = number of cached keys before it)
 *   out: hidden bf16 [n_seq, C] = ln_f output; x; the k / v rows of the new token in every layer's cache at (row, position)
 * dh_decode_plan_build encodes the TMA descriptors and the stage table into a HOST blob of dh_decode_plan_bytes(n_layer) bytes;
 * the caller copies it to device memory (64-byte aligned) and passes both copies to dh_decode_layers.  `barrier` = 16 bytes of
 * device memory zeroed once; `ws` >= dh_decode_workspace_bytes(desc).  Cache row of sequence i is i.  ABI v3.
 * ------------------------------------------------------------------------------------------------- */
typedef struct dh_decode_layer {
  const void* wqkv; const void* wproj; const void* w12; const void* wmlp;   /* bf16 [N, K] as stored; w12 = fc_1|fc_2 in 32-row blocks */
  const float* norm2;              /* this layer's norm_2 weight */
  const float* norm_next;          /* the next layer's norm_1 weight; ln_f weight for the last layer */
  void* cache_k; void* cache_v;    /* bf16 [n_seq, n_groups, max_len, 64] */
  const void* prefix_kv;           /* adapter prefix rows bf16 [prefix_len, n_groups*(qpk+2)*64] or NULL */
  const float* gate;               /* f32 [n_head] or NULL */
} dh_decode_layer;

typedef struct dh_decode_desc {
  int32_t n_layer, n_seq, C, I, n_groups, q_per_kv, max_len, prefix_len;
  float eps;
  void* x; void* x2; void* n; void* qkv; void* y; void* act; void* hidden;
  void* ws; size_t ws_bytes;
  const int32_t* positions;
  const void* rope_cos; const void* rope_sin;
  void* barrier;
  const dh_decode_layer* layers;
  void* trace;                     /* developer trace, int64 [3 * n_layer * 24 + 8 * n_sm * 2] globaltimer stamps (NULL: off); see scripts/decode_mega_trace.py */
} dh_decode_desc;

size_t dh_decode_plan_bytes(int n_layer);
size_t dh_decode_workspace_bytes(const dh_decode_desc* desc);
int dh_decode_plan_build(const dh_decode_desc* desc, void* plan_host);
int dh_decode_layers(const void* plan_dev, const void* plan_host, void* stream);

/* Scatter the k and v slots of qkv rows into the cache at (seq[m], pos[m]). */
int dh_kv_cache_write(const void* qkv, void* cache_k, void* cache_v, const int32_t* seq, const int32_t* pos, int M,
                      int max_len, int n_groups, int q_per_kv, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * Loss (ger/utils.py:424-463 over lm_head logits, ger/lora.py:546-549)
 * logits bf16 [M, V] (ld = ldl).  targets int64 [M], -1 = ignore.  Writes per-row loss f32 [M] (0 where ignored)
 * and, when write_grad, overwrites logits IN PLACE with d(loss)/d(logits) * grad_scale
 *   = (softmax - onehot) * grad_scale   for valid rows, 0 for ignored rows (write_grad = 2: ignored rows are left untouched).
 * The caller owns the reduction: train = sum(row_loss)/M (quirk Q2), val = sum/valid.
 * ------------------------------------------------------------------------------------------------- */
int dh_cross_entropy(void* logits, int64_t ldl, const int64_t* targets, float* row_loss, int M, int V,
                     float grad_scale, int write_grad, void* stream);
/* Valid-row compaction for the fused lm_head + loss (only rows with a target != -1 carry loss or gradient, so only they
 * are projected onto the vocabulary).  count int32[1] = number of valid targets; rows int32[M]: rows[i] = row of the i-th
 * valid target (ascending), -1 for i >= count; targets_c int64[M] = the gathered targets, -1 beyond count.
 * dh_gather_rows: dst[i] = src[rows[i]] for i < *count (bf16 [., C]).  dh_scatter_rows: dst [M, C] = 0 except
 * dst[rows[i]] = src[i].  With dh_gemm's m_limit = count and dh_cross_entropy's write_grad = 2 nothing beyond the
 * valid rows (rounded up to one GEMM tile) is computed, without the host ever learning the count. */
int dh_compact_targets(const int64_t* targets, int M, int32_t* count, int32_t* rows, int64_t* targets_c, void* stream);
int dh_gather_rows(const void* src, const int32_t* rows, const int32_t* count, void* dst, int M, int C, void* stream);
int dh_scatter_rows(const void* src, const int32_t* rows, const int32_t* count, void* dst, int M, int C, void* stream);
/* out[0] = sum(row_loss[0..M)), out[1] = number of targets != -1 */
int dh_loss_reduce(const float* row_loss, const int64_t* targets, float* out2, int M, void* stream);
/* argmax over bf16 logits rows (generate/base.py:62-70 with top_k=1): ids int64 [M]; ties -> lowest index */
int dh_argmax(const void* logits, int64_t ldl, int64_t* ids, int M, int V, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * RelPrompt reliability predictor (ger/relprompt.py:126-147): Conv1d(k3,p1)+ReLU -> Conv1d(k3,p1)+ReLU ->
 * AvgPool1d(pool, ceil_mode) -> Linear(256->3).  x f32 [B, T, C]; conv weights f32 in PyTorch layout
 * [256, Cin, 3]; logits f32 [B, ceil(T/pool), 3].  h1/h2 f32 [B, T, 256] are scratch that the backward reads.
 * ------------------------------------------------------------------------------------------------- */
int dh_relpred_fwd(const float* x, const float* w1, const float* b1, const float* w2, const float* b2,
                   const float* wc, const float* bc, float* h1, float* h2, float* pooled, float* logits, int B, int T,
                   int C, int pool, void* stream);
/* The tail of the head alone: AvgPool1d(pool, ceil_mode) + Linear(256->3) over h2 = relu(conv2) f32 [B, T, 256], and its backward
 * (dz2 f32 [B, T, 256] = gradient of conv2's pre-activation; dwc / dbc ACCUMULATED).  The default (tensor-core) path of
 * `NoiseMaskClassifier` runs the two convolutions as bf16 GEMMs through dh_gemm -- a row of three consecutive frames is the
 * K-major A operand -- and calls these for the rest; dh_relpred_fwd / dh_relpred_bwd below stay as the fp32 path
 * (DH_RELPRED=fp32, the arithmetic of the reference on CPU). */
int dh_relpred_head_fwd(const float* h2, const float* wc, const float* bc, float* pooled, float* logits, int B, int T, int pool,
                        void* stream);
int dh_relpred_head_bwd(const float* h2, const float* pooled, const float* d_logits, const float* wc, float* dz2, float* dwc,
                        float* dbc, int B, int T, int pool, void* stream);
/* d_logits f32 [B, n_out, 3] -> gradients of all six parameter tensors (ACCUMULATED; caller zeroes). */
int dh_relpred_bwd(const float* x, const float* w1, const float* w2, const float* wc, const float* h1, const float* h2,
                   const float* pooled, const float* d_logits, float* dw1, float* db1, float* dw2, float* db2,
                   float* dwc, float* dbc, float* scratch, int B, int T, int C, int pool, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * Small utilities
 * ------------------------------------------------------------------------------------------------- */
/* dst bf16 [rows, cols] = src f32 [rows, cols] (ld in elements) */
int dh_cast_f32_to_bf16(const float* src, int64_t lds, void* dst, int64_t ldd, int64_t rows, int64_t cols, void* stream);
/* nn.Dropout(p) on a bf16 tensor of n elements (the input of the LoRA branch, ger/lora.py:94-98,165; finetune/ger.py:401 trains with
 * p = 0.05): out = keep ? x / (1 - p) : 0.  The mask is Philox4x32-10 of (seed, offset, element index / 8): the same (seed, offset)
 * regenerates it, so the backward recomputes dropout(x) for the factor gradients instead of saving it.  accumulate != 0: out += the
 * masked, scaled x (the LoRA branch's gradient flowing back into x through the mask).  The keep probability is quantised to
 * 1 - round(p * 65536) / 65536. */
int dh_dropout_bf16(const void* x, void* out, int64_t n, float p, uint64_t seed, uint64_t offset, int accumulate, void* stream);
/* fused AdamW over one flat fp32 bucket (torch.optim.AdamW semantics, finetune/ger.py:132) */
int dh_adamw_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr, float beta1,
                  float beta2, float eps, float weight_decay, int step, float grad_scale, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DUALHYP_H */
