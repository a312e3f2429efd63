/*
 * dli_b200.h — C ABI of the B200-native span forward (libdli_b200.so).
 *
 * The reference (Dylan102938/distributed-llm-inference) is pure Python: it has NO FFI / plugin boundary.
 * The boundary this library sits behind is the Python interface of its hot path:
 *   LlamaBlock.forward                     distributed_llm_inference/models/llama/model.py:25-76
 *   OptimizedLlamaDecoderLayer.forward     distributed_llm_inference/models/llama/modules.py:146-184
 *   OptimizedLlamaInferenceAttention.fwd   distributed_llm_inference/models/llama/modules.py:36-114
 *   PartialLlamaSinkCache.update           distributed_llm_inference/models/llama/cache.py:64-135
 * Each entry point below names the reference lines it replaces; the ctypes stub a maintainer of the
 * reference would add is in INTEGRATION.md and shipped as distributed_llm_inference_b200/_ffi.py.
 *
 * Conventions
 *   - plain C types only; every function returns an int status (0 = DLI_OK, negative = error) and never
 *     throws; dli_last_error() returns a thread-local message for the last failure on this thread.
 *   - every buffer is CALLER-ALLOCATED device memory (in practice PyTorch's allocator) and is never freed
 *     or retained by the library beyond the call, except weights bound with dli_span_bind_weights, which
 *     are borrowed until the span is destroyed.  Unless an entry point's comment says otherwise the
 *     library allocates no device memory.
 *   - explicit stream argument (a cudaStream_t passed as void*); no hidden synchronisation; the forward
 *     path does no host allocation.  Per-call state that changes between decode steps (sequence lengths,
 *     block tables, cos/sin rows) is read from DEVICE memory so that a captured CUDA graph can be replayed.
 *   - calls that share a workspace must be serialised by the caller (the reference's runtime calls one
 *     pool function at a time per device).
 *   - activations and weights are bf16 (DLI_BF16).  Weights are row-major [out_features, in_features]
 *     exactly as nn.Linear stores them.
 */
#ifndef DLI_B200_H
#define DLI_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DLI_VERSION 1

enum dli_status {
  DLI_OK = 0,
  DLI_ERR_ARG = -1,          /* bad argument (null pointer, shape not supported, misalignment) */
  DLI_ERR_CUDA = -2,         /* a CUDA runtime / driver call failed; message has the CUDA error string */
  DLI_ERR_UNSUPPORTED = -3,  /* valid request outside what the sm_100a kernels implement */
  DLI_ERR_WORKSPACE = -4     /* workspace too small */
};

/* Activations and weights are bf16, the dtype BASELINE's north star names.  The reference casts checkpoints to fp16
 * (utils/model.py:67-68); tcgen05 kind::f16 multiplies fp16 as well, but no fp16 instantiation exists yet: any other
 * value is refused by dli_span_create. */
enum dli_dtype { DLI_BF16 = 0 };

typedef void* dli_stream_t; /* cudaStream_t */

int dli_version(void);
const char* dli_last_error(void);

/* ------------------------------------------------------------------------------------------------
 * Paged KV cache layout (replaces the per-generation torch.cat lists of cache.py:14-15,103-109).
 * One pool per layer: K pages and V pages are arrays [n_pages][n_kv_heads][page_tokens][head_dim] bf16.
 * Keys are stored ROTATED (post-RoPE), as the reference stores them (modules.py:76-85).
 * block_table[b * max_pages + j] = physical page holding tokens [j*page_tokens, (j+1)*page_tokens) of
 * sequence b.  seq_lens[b] = number of tokens stored for sequence b BEFORE this call.
 * ---------------------------------------------------------------------------------------------- */
typedef struct dli_kv_desc {
  void* k_pages;              /* device, layer 0 */
  void* v_pages;              /* device, layer 0 */
  int64_t layer_stride_bytes; /* byte distance between consecutive layers' pools */
  int32_t page_tokens;        /* 16, 32 or 64 */
  int32_t n_pages;            /* pages per layer pool */
  const int32_t* block_table; /* device [batch, max_pages] */
  const int32_t* seq_lens;    /* device [batch]: tokens cached before this call */
  int32_t max_pages;          /* row length of block_table */
  int32_t max_ctx;            /* host-side upper bound on seq_lens[b] + q_len for every b (sizes the grid) */
} dli_kv_desc;

/* ------------------------------------------------------------------------------------------------
 * Operator-level entry points (one per fused kernel).  Used by the parity tests and by callers that
 * want to assemble their own layer; dli_span_forward chains the same kernels.
 * ---------------------------------------------------------------------------------------------- */

/* RMSNorm: y = bf16( w * (s*x) * rsqrt(mean((s*x)^2) + eps) ), fp32 math.  x is fp32 or bf16 [rows, h].
 * If resid_out != NULL the fp32 copy of x is also written there (entry of a span: bf16 hand-off ->
 * fp32 residual stream).
 * Replaces LlamaRMSNorm.forward as called at modules.py:162 (s = 1) and modules.py:179 (s = 2: the
 * reference normalises hidden_states + residual, i.e. 2*h). eps default there is 1e-6 (modules.py:124-125). */
int dli_rmsnorm(const void* x, int x_is_f32, const void* weight_bf16, void* y_bf16, float* resid_out,
                int rows, int h, float eps, float pre_scale, dli_stream_t stream);

/* GEMM epilogues */
enum dli_gemm_epilogue {
  DLI_EPI_STORE_F32 = 0,  /* out_f32[m, n]  = acc                                   */
  DLI_EPI_ADD_F32 = 1,    /* out_f32[m, n] += acc   (fp32 reduction in L2; split-K capable)  */
  DLI_EPI_STORE_BF16 = 2, /* out_bf16[m, n] = acc                                   */
  DLI_EPI_SILU_MUL_BF16 = 3 /* w = [gate; up] ([2*n, k]); out_bf16[m, n] = silu(x.gate_n) * (x.up_n) */
};

/* out[m, n] (+)= sum_k x[m, k] * w[n, k]  — tcgen05/TMEM GEMM fed by TMA, fp32 accumulation.
 * x bf16 [m, k] row-major (ldx elements), w bf16 [n(*2), k] row-major (nn.Linear layout), out [m, n] (ldo).
 * k % 8 == 0, ldx % 8 == 0; pointers 16-byte aligned.  m <= 128 takes the weight-streaming (swap-AB,
 * split-K) path used at decode; larger m the 128x128-tile path used at prefill.
 * Replaces the seven nn.Linear calls of a layer: q/k/v_proj modules.py:61-63, o_proj modules.py:112,
 * LlamaMLP gate/up/down modules.py:181 — with the residual adds of modules.py:173,182 (ADD_F32) and
 * silu(gate)*up (SILU_MUL_BF16) folded into the epilogue. */
int dli_gemm(const void* x_bf16, int64_t ldx, const void* w_bf16, void* out, int64_t ldo, int m, int n,
             int k, int epilogue, dli_stream_t stream);

/* RoPE + KV append.  qkv_f32 [tokens, (H + 2*KV) * hd] is the fused q|k|v projection output, token
 * t = b * q_len + l.  cos/sin are fp32 [q_len, hd] (one position vector shared by the batch, as in the
 * reference: modules.py:17-20 broadcasts cos [1, L, hd]).  Writes rotated q (bf16 [tokens, H*hd]) and
 * appends rotated k and v at positions seq_lens[b] + l of sequence b's pages of `layer`.  zero_after != 0
 * clears qkv_f32 after reading it (it is a split-K accumulation buffer on the decode path).
 * Replaces apply_rotary_pos_emb modules.py:17-20,73-76 and the append branch of
 * PartialLlamaSinkCache.update cache.py:103-109. */
int dli_rope_kv_append(float* qkv_f32, int zero_after, const float* cos, const float* sin, void* q_bf16,
                       const dli_kv_desc* kv, int layer, int batch, int q_len, int n_heads,
                       int n_kv_heads, int head_dim, dli_stream_t stream);

/* Scratch floats needed by dli_decode_attention for the given problem (split-K partials). */
size_t dli_decode_attention_scratch_bytes(int batch, int n_heads, int head_dim, int max_ctx);

/* Decode attention (q_len == 1): out[b, h, :] = softmax(q[b,h,:] . K[b, h/G, 0:T_b, :]^T / sqrt(hd)) . V,
 * T_b = seq_lens[b] + 1 (the new token must already be appended).  Split-K flash-decode over the
 * paged cache with fp32 online softmax; GQA by indexing (no repeat_kv copy).  scratch: fp32, at least
 * dli_decode_attention_scratch_bytes(...) bytes, caller-allocated.
 * Replaces repeat_kv + QK^T + softmax + PV + head merge, modules.py:87-105, for q_len == 1. */
int dli_decode_attention(const void* q_bf16, void* out_bf16, const dli_kv_desc* kv, int layer,
                         int batch, int n_heads, int n_kv_heads, int head_dim, float* scratch,
                         size_t scratch_bytes, dli_stream_t stream);

/* Decode attention for callers that cannot allocate.  Same computation and arguments as dli_decode_attention,
 * minus the scratch: the split-K partials go to an 8 MiB array that is built into the library (each GPU that
 * loads libdli_b200.so holds those 8 MiB until the process exits).  Meant for code that receives only its input
 * and output buffers — e.g. an operator plug-in inside a host framework that has reserved all of HBM for its own
 * pool, or a serving loop that cannot afford cudaMalloc's device-wide synchronisation or its failure mid-request.
 * It never calls cudaMalloc, cudaFree or any allocating library routine.
 *   - Limit: batch * n_heads * splits * (head_dim + 2) * 4 bytes must fit in 8 MiB (splits <= 32 is chosen from
 *     batch, n_kv_heads and kv->max_ctx as in dli_decode_attention); larger problems are refused with
 *     DLI_ERR_WORKSPACE and a message naming both sizes.
 *   - The first call on each device looks up the array's device address (cudaGetSymbolAddress, with that device
 *     current) and remembers it; make that first call before any stream capture begins.
 *   - All calls on a device share the one array.  Nothing in it has to be zero beforehand (every slot a call
 *     reads it has written itself), but two calls that overlap on the GPU — different streams, or graphs
 *     replayed concurrently — race on it: serialise them.
 * This entry point has no counterpart in the reference. */
int dli_decode_attention_static(const void* q_bf16, void* out_bf16, const dli_kv_desc* kv, int layer,
                                int batch, int n_heads, int n_kv_heads, int head_dim, dli_stream_t stream);

/* Prefill attention (q_len >= 1, with or without cached past): queries of a chunk at positions
 * seq_lens[b] .. seq_lens[b]+q_len-1 attend to keys [0, seq_lens[b] + q_len) (causal == 0: the
 * reference's default sdpa configuration adds no mask, model.py:94-101) or to keys
 * [0, seq_lens[b] + l] (causal == 1: eager mask of model.py:103-127).  The chunk's k/v must already be
 * appended.  Tensor-core flash attention, fp32 online softmax.  Replaces modules.py:87-105. */
int dli_prefill_attention(const void* q_bf16, void* out_bf16, const dli_kv_desc* kv, int layer,
                          int batch, int q_len, int n_heads, int n_kv_heads, int head_dim, int causal,
                          dli_stream_t stream);

/* Masks beyond the causal flag (LlamaBlock._update_causal_mask, model.py:103-135):
 *   key_limit  device int32 [q_len] or NULL: query l sees cache slots j < key_limit[l].  Replaces the causal rule; needed
 *              when a multi-token chunk crosses the attention-sink window, where the reference compares the ABSOLUTE
 *              cache_position of a query with the SLOT index of a key in the evicted cache (model.py:110 after
 *              cache.py:111-133), i.e. key_limit[l] = min(window, cache_position[l] + 1).
 *   key_valid  device uint8 [batch, key_stride] or NULL: 0 marks a padded key (attention_mask == 0, model.py:128-135),
 *              visible to no query of that sequence; key_stride >= kv->max_ctx.
 * Hidden scores take a large finite negative value like the reference's additive finfo.min, so a prefill row with no
 * visible key (a left-padding query) is the uniform average of V over the context, as in the reference.  At decode a
 * hidden key has probability 0 (the step's own key is always valid).  These calls take the generic kernels (mma.sync
 * prefill, register-staged decode): padded batches and window-crossing chunks are the rare path. */
typedef struct dli_attn_mask {
  const int32_t* key_limit;
  const uint8_t* key_valid;
  int32_t key_stride;
} dli_attn_mask;

int dli_prefill_attention_masked(const void* q_bf16, void* out_bf16, const dli_kv_desc* kv, int layer, int batch,
                                 int q_len, int n_heads, int n_kv_heads, int head_dim, int causal,
                                 const dli_attn_mask* mask, dli_stream_t stream);
int dli_decode_attention_masked(const void* q_bf16, void* out_bf16, const dli_kv_desc* kv, int layer, int batch,
                                int n_heads, int n_kv_heads, int head_dim, const dli_attn_mask* mask, float* scratch,
                                size_t scratch_bytes, dli_stream_t stream);

/* Attention-sink eviction for every layer of a span, in place on the pages (call BEFORE the forward that appends
 * the n_new tokens, when cur_len + n_new >= window): the cache becomes [first `sinks` tokens | last
 * (window - sinks - n_new) tokens with keys re-rotated back by n_new positions]; afterwards the caller sets
 * seq_lens[b] = window - n_new.  cos_hist/sin_hist: fp32 [window, hd], the cos/sin rows of the generation's first
 * `window` positions (the reference's _cos_cache/_sin_cache).  All sequences of the generation share cur_len.
 * Replaces the eviction branch of PartialLlamaSinkCache.update, cache.py:111-133, and
 * _get_rerotation_cos_sin, cache.py:21-48. */
int dli_kv_evict(const dli_kv_desc* kv, int n_layers, int batch, int n_kv_heads, int head_dim, int cur_len,
                 int window, int sinks, int n_new, const float* cos_hist, const float* sin_hist,
                 dli_stream_t stream);

/* act_bf16[m, n] = silu(gu_f32[m, n]) * gu_f32[m, n_total + n]; gu is [m, 2*n].  Decode-path companion of
 * the split-K gate|up GEMM (the activation of LlamaMLP.forward, modules.py:181). If zero_after != 0 the
 * fp32 input is cleared after it is read (it is a split-K accumulation buffer). */
int dli_silu_mul(float* gu_f32, void* act_bf16, int m, int n, int zero_after, dli_stream_t stream);

/* out_bf16 = bf16(in_f32), n elements (span exit: fp32 residual stream -> bf16 hand-off tensor). */
int dli_cast_f32_bf16(const float* in, void* out_bf16, int64_t n, dli_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Span-level entry point: the whole LlamaBlock.forward loop (model.py:59-71) in one call.
 * ---------------------------------------------------------------------------------------------- */
typedef struct dli_span dli_span; /* host-side object: shapes, borrowed weight pointers, TMA descriptors */

typedef struct dli_span_desc {
  int32_t hidden_size, intermediate_size, n_heads, n_kv_heads, head_dim, n_layers;
  int32_t dtype;        /* DLI_BF16 */
  float eps_input;      /* 1e-6: LlamaRMSNorm(hidden_size) default, modules.py:124 */
  float eps_post;       /* 1e-6, modules.py:125 */
  float post_prescale;  /* 2.0: post_attention_layernorm(hidden_states + residual), modules.py:179 */
} dli_span_desc;

typedef struct dli_layer_weights {
  const void* wqkv;   /* [(H + 2*KV) * hd, h]  rows: q_proj | k_proj | v_proj   */
  const void* wo;     /* [h, H * hd]                                            */
  const void* wgu;    /* [2 * i, h]             rows: gate_proj | up_proj       */
  const void* wdown;  /* [h, i]                                                 */
  const void* g_in;   /* [h] input_layernorm.weight                             */
  const void* g_post; /* [h] post_attention_layernorm.weight                    */
} dli_layer_weights;

int dli_span_create(const dli_span_desc* desc, dli_span** out);
int dli_span_destroy(dli_span* span);
int dli_span_bind_weights(dli_span* span, int layer, const dli_layer_weights* w);

/* ------------------------------------------------------------------------------------------------
 * Weight-only int8 (the weight format of the reference's optimised mode: convert_to_optimized_block turns every
 * nn.Linear into bitsandbytes Linear8bitLt(has_fp16_weights=False), utils/model.py:93-123, i.e. row-wise absmax int8
 * with one fp32 scale per output row).  q[r, k] = round(w[r, k] / scale[r]) + 128 as uint8, scale[r] = max|w[r, :]| / 127.
 * Activations stay bf16: the decode GEMM widens the int8 tile to bf16 in shared memory (exact) and scales the fp32
 * accumulator by scale[r] -- half the weight bytes per decode step, no activation quantisation, no outlier columns.
 * Chunks of more than 128 tokens (prefill) dequantise one matrix at a time into a caller-allocated bf16 scratch and
 * run the bf16 GEMMs on it.
 * ---------------------------------------------------------------------------------------------- */
int dli_quantize_rows_w8(const void* w_bf16, int64_t n, int64_t k, void* q_u8, float* scale, dli_stream_t stream);
int dli_dequantize_rows_w8(const void* q_u8, const float* scale, int64_t n, int64_t k, void* w_bf16, dli_stream_t stream);
/* out[m, n] (+)= scale[n] * sum_k x[m, k] * (q[n, k] - 128), m <= 128, k % 16 == 0; epilogues as dli_gemm (for
 * DLI_EPI_SILU_MUL_BF16, q and scale hold 2n rows: gate then up). */
int dli_gemm_w8(const void* x_bf16, int64_t ldx, const void* q_u8, const float* scale, void* out, int64_t ldo, int m,
                int n, int k, int epilogue, dli_stream_t stream);

typedef struct dli_layer_weights_w8 {
  const void* wqkv;  const float* sqkv;   /* uint8 [(H + 2*KV) * hd, h], fp32 [(H + 2*KV) * hd] */
  const void* wo;    const float* so;     /* [h, H * hd], [h]                                    */
  const void* wgu;   const float* sgu;    /* [2 * i, h], [2 * i]                                 */
  const void* wdown; const float* sdown;  /* [h, i], [h]                                         */
  const void* g_in;                       /* bf16 [h] input_layernorm.weight                     */
  const void* g_post;                     /* bf16 [h] post_attention_layernorm.weight            */
} dli_layer_weights_w8;
int dli_span_bind_weights_w8(dli_span* span, int layer, const dli_layer_weights_w8* w);
/* bf16 scratch for the prefill path of a span with int8 weights: at least dli_span_dequant_scratch_bytes(span) bytes,
 * 256-byte aligned, caller-allocated, borrowed until replaced or the span is destroyed. */
size_t dli_span_dequant_scratch_bytes(const dli_span* span);
int dli_span_set_dequant_scratch(dli_span* span, void* scratch_bf16, size_t bytes);

/* ------------------------------------------------------------------------------------------------
 * int8 activations on int8 weights -- the arithmetic of the reference's optimised mode itself: bitsandbytes
 * Linear8bitLt(threshold, has_fp16_weights=False), utils/model.py:93-123 ("LLM.int8"): per token row the entries below
 * the threshold are quantised with the row's absmax over those entries (scale = absmax / 127, code = round(x / scale)),
 * the product runs int8 x int8 -> int32 on the tensor cores (tcgen05.mma.kind::i8) and is scaled by x_scale[m] *
 * w_scale[n]; the entries at or above the threshold (outliers) do not enter that scale.  bitsandbytes multiplies their
 * columns in fp16; here they form a SECOND int8 plane of the same row with its own scale (absmax over the outliers /
 * 127), multiplied against the same weight tiles into a second accumulator -- the outliers are kept to 1/254 of the
 * row's largest one, at a cost independent of their number.  Decode shapes only (m <= 128).
 * A dli_act_i8 describes one quantised activation matrix [m, k] in caller-allocated device memory:
 *   q int8 [m, k] (16-byte aligned, k % 16 == 0), scale fp32 [m], ksum int32 [m, ceil(k / 128)] (sum of the row's codes
 *   per 128-element block: the weights are stored biased, so the GEMM subtracts 128 * ksum); q2 / scale2 / ksum2 the
 *   same for the outlier plane, or all NULL for a one-plane activation (no threshold: plain row-wise absmax).
 * ---------------------------------------------------------------------------------------------- */
typedef struct dli_act_i8 {
  void* q;
  float* scale;
  int32_t* ksum;
  void* q2;
  float* scale2;
  int32_t* ksum2;
} dli_act_i8;
/* bytes of one dli_act_i8 with `planes` (1 or 2) planes for up to m rows of k entries (each array 256-byte aligned),
 * and its carving */
size_t dli_act_i8_bytes(int m, int k, int planes);
int dli_act_i8_carve(void* base, size_t bytes, int m, int k, int planes, dli_act_i8* out);
/* threshold <= 0 or a one-plane destination: every entry goes to the first plane */
int dli_quantize_act_i8(const void* x_bf16, int64_t ldx, int m, int k, float threshold, const dli_act_i8* dst,
                        dli_stream_t stream);
/* out[m, n] (+)= w_scale[n] * sum over planes of scale_p[m] * sum_k q_p[m, k] * (q[n, k] - 128), n % 4 == 0; epilogues
 * as dli_gemm_w8 (DLI_EPI_SILU_MUL_BF16: q and w_scale hold 2n rows, gate then up) */
int dli_gemm_i8(const dli_act_i8* x, const void* q_u8, const float* w_scale, void* out, int64_t ldo, int m, int n, int k,
                int epilogue, dli_stream_t stream);
/* A span bound with dli_span_bind_weights_w8 runs its decode GEMMs (chunks of <= 128 tokens) in this mode once it has
 * been given a scratch of dli_span_act_i8_scratch_bytes(span) bytes (256-byte aligned, caller-allocated, borrowed until
 * replaced or the span is destroyed; NULL switches back to bf16 activations).  threshold <= 0: one plane, no outliers. */
size_t dli_span_act_i8_scratch_bytes(const dli_span* span);
int dli_span_set_act_i8(dli_span* span, void* scratch, size_t bytes, float threshold);

/* Bytes of caller-allocated device workspace dli_span_forward needs for chunks of up to max_tokens
 * (= batch * q_len) tokens, batch <= max_batch, context <= max_ctx. */
size_t dli_span_workspace_bytes(const dli_span* span, int max_tokens, int max_batch, int max_ctx);

/* hidden_in / hidden_out: bf16 [batch * q_len, h] (may alias).  cos/sin fp32 [q_len, hd].
 * Appends q_len tokens per sequence to the cache (the caller advances seq_lens afterwards) and runs
 * every bound layer: norm -> qkv -> rope/append -> attention -> o_proj(+res) -> norm(2h) -> gate|up
 * -> silu*up -> down(+res).  all_hidden (nullable): bf16 [n_layers + 1, tokens, h] receives the input of
 * every layer and the final output (output_hidden_states, model.py:57-61,73-74).
 * The workspace must be zero-filled ONCE after allocation: the split-K accumulation buffers at its
 * front are re-zeroed by the kernels that consume them and must read zero at every entry (so a call
 * that failed half-way requires a fresh memset).
 * Replaces LlamaBlock.forward model.py:25-76 (mask construction model.py:78-144 is reduced to the
 * `causal` flag; cos/sin of model.py:55 are inputs so that the reference's int64-truncated tables and
 * the intended float tables are both host-side choices). */
int dli_span_forward(dli_span* span, const dli_kv_desc* kv, int batch, int q_len, int causal,
                     const void* hidden_in, void* hidden_out, const float* cos, const float* sin,
                     void* all_hidden, void* workspace, size_t workspace_bytes, dli_stream_t stream);

/* dli_span_forward with a mask beyond the causal flag (see dli_attn_mask): padded batches (attention_mask with zeros,
 * model.py:128-135) and multi-token chunks that cross the sink window (cache.py:107-133 with n > 1). */
int dli_span_forward_masked(dli_span* span, const dli_kv_desc* kv, int batch, int q_len, int causal,
                            const dli_attn_mask* mask, const void* hidden_in, void* hidden_out, const float* cos,
                            const float* sin, void* all_hidden, void* workspace, size_t workspace_bytes,
                            dli_stream_t stream);

/* One decode step (q_len == 1) whose bookkeeping lives on the device, for replay in a CUDA graph: the same kernels
 * as dli_span_forward(..., q_len = 1, ...), but
 *   - cos_table / sin_table are the WHOLE fp32 [max_positions, head_dim] tables and the kernels use row *position
 *     (device int64) -- no gather launch in front of the step;
 *   - advance != 0: the span's last kernel adds 1 to kv->seq_lens[b] for b < batch and to *position, so the next
 *     replay sees the next token -- no length/position update launches behind the step.
 * The caller keeps *position below max_positions (the host knows how many steps it replayed).
 * Replaces, at decode, the per-call position/rotary set-up of LlamaBlock.forward (model.py:44-55) and the length
 * bookkeeping of PartialLlamaSinkCache.update (cache.py:86-87,103-109). */
int dli_span_decode_step(dli_span* span, const dli_kv_desc* kv, int batch, const void* hidden_in, void* hidden_out,
                         const float* cos_table, const float* sin_table, long long* position, int advance,
                         void* workspace, size_t workspace_bytes, dli_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Pipeline stage step: dli_span_decode_step plus the hand-off between consecutive spans, which in the reference is
 * a network hop through hivemind (only imported, server/backend.py:4-7; SURVEY 8b "dli_p2p_*", 8e) and here is done by
 * the step's own first and last kernels over NVLink peer memory:
 *   - INPUT   hidden_in: bf16 [batch, h] in THIS GPU's memory (a hand-off inbox the previous stage stores into), or
 *             token_ids + embed_table: device int32 [batch] ids whose rows of the bf16 [vocab, h] table are the input
 *             (stage 0 of a ring that carries tokens: the embedding the client would look up).
 *   - WAIT    wait_flag (this GPU's memory, nullable): the entry kernel spins on it (ld.acquire.sys) until
 *             *wait_flag + wait_lag >= state[0] + 1, i.e. until the producer has finished the step that feeds this one.
 *             wait_lag = 0 when the producer runs the same round earlier (stage r-1 -> r), 1 when it is a round behind
 *             (ring closure: last stage -> stage 0, which therefore starts round 0 without waiting).
 *   - OUTPUT  hidden_out: bf16 [batch, h], this GPU's or a PEER's memory (the next stage's inbox), nullable when the
 *             stage ends in the head; lm_head (nullable): bf16 [vocab, h] -- final RMSNorm (final_norm, eps_final) ->
 *             logits fp32 [batch, vocab] (this GPU) -> greedy argmax -> tokens_out int32 [batch] (this GPU's or a
 *             peer's memory: stage 0's token inbox).  The tail of the client (norm, head, argmax; absent from the
 *             reference, model.py:19-23) as three more kernels of the same chain.
 *   - SIGNAL  after all of the last kernel's stores are fenced system-wide, state[0] += 1 and, if signal_flag is not
 *             NULL, *signal_flag = state[0] (st.release.sys; usually the consumer's wait_flag in peer memory).
 *   state: 4 x uint32 in this GPU's memory, zero-initialised by the caller: [0] steps completed, [1] ticket (0 between
 *   steps), [2] sticky error: a wait gave up after timeout_ms (default 2000) -- later waits return at once and the step
 *   computes garbage instead of hanging the GPU; the host checks it after synchronising.  [3] unused.
 *   chain != 0: the entry kernel is launched with programmatic dependent launch like the kernels inside the step, no
 *   stream barrier in front of it.  Only valid when the launches that precede the step in the stream belong to OTHER
 *   generations (the other micro-batches of a pipeline round, a whole round between two steps of the same one):
 *   pre-wait prologues read the generation's lengths and past K/V.
 * Everything else (cos/sin tables, *position, advance, workspace) is as in dli_span_decode_step. */
typedef struct dli_stage_io {
  const void* hidden_in;
  const int32_t* token_ids;
  const void* embed_table;
  void* hidden_out;
  const void* final_norm; /* bf16 [h] */
  const void* lm_head;    /* bf16 [vocab, h] */
  float* logits;
  int32_t* tokens_out;
  int32_t vocab;
  float eps_final;
  const uint32_t* wait_flag;
  uint32_t* signal_flag;
  uint32_t* state;
  int32_t wait_lag;
  int32_t chain;
  int32_t timeout_ms;
} dli_stage_io;

int dli_stage_decode_step(dli_span* span, const dli_kv_desc* kv, int batch, const dli_stage_io* io,
                          const float* cos_table, const float* sin_table, long long* position, int advance,
                          void* workspace, size_t workspace_bytes, dli_stream_t stream);

/* Stream-ordered wait (one thread) until *flag has reached at_least (counters compared modulo 2^32), with the same
 * time-out and sticky error word (state[2]) as the in-kernel wait: what a client puts on a side stream in front of
 * the copy that reads a result another GPU delivers.  No counterpart in the reference. */
int dli_flag_wait(const uint32_t* flag, uint32_t at_least, uint32_t* state, int timeout_ms, dli_stream_t stream);

/* Client tail as one call: xn = RMSNorm(hidden; final_norm, eps) (bf16 [batch, h] scratch), logits = xn . lm_head^T
 * (fp32 [batch, vocab]), tokens_out[b] = argmax_v logits[b, v] (int32, lowest index on ties).  hidden is bf16 or fp32
 * [batch, h], batch <= 128.  The reference builds neither norm nor head (model.py:19-23); this is what its client
 * would run after the last span. */
int dli_lm_head_argmax(const void* hidden, int hidden_is_f32, const void* final_norm_bf16, float eps,
                       const void* lm_head_bf16, int vocab, int h, int batch, void* xn_bf16, float* logits,
                       int32_t* tokens_out, dli_stream_t stream);

/* Per-kernel-class timing of dli_span_forward, measured with CUDA events recorded on the call's own stream
 * at every class boundary (bench.py's roofline figures).  dli_span_profile(span, 1) starts collecting (host-side
 * cudaEvent objects are created on demand; no device memory); every forward after that appends intervals;
 * dli_span_profile_read synchronises on the last event, returns the summed milliseconds and the number of
 * intervals per class since the last read, and resets.  Profiling adds an event record between kernels and is
 * off by default.  No counterpart in the reference (it has no instrumentation). */
enum dli_prof_class {
  DLI_PROF_RMSNORM = 0, DLI_PROF_GEMM_QKV = 1, DLI_PROF_ROPE_APPEND = 2, DLI_PROF_ATTENTION = 3,
  DLI_PROF_GEMM_O = 4, DLI_PROF_GEMM_GATE_UP = 5, DLI_PROF_SILU_MUL = 6, DLI_PROF_GEMM_DOWN = 7,
  DLI_PROF_CAST = 8, DLI_PROF_CLASSES = 9
};
int dli_span_profile(dli_span* span, int enable);
int dli_span_profile_read(dli_span* span, float* ms_by_class, int* intervals_by_class);

/* Number of kernel launches (incl. memset nodes) the last dli_span_forward on this thread issued. */
int dli_last_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* DLI_B200_H */
