// extern "C" surface of libdli_b200 and the span runner (LlamaBlock.forward, model.py:25-76, as one call).
#include "common.cuh"

#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <new>
#include <vector>

namespace dli {

// ---- error / bookkeeping --------------------------------------------------------------------------
static thread_local char g_err[512] = "";
static thread_local int g_launches = 0;

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
void count_launch(int n) { g_launches += n; }
void reset_launch_count() { g_launches = 0; }

int current_device(int* dev) {
  DLI_CUDA_OK(cudaGetDevice(dev));
  if (*dev < 0 || *dev >= DLI_MAX_DEVICES) return fail(DLI_ERR_UNSUPPORTED, "device ordinal %d outside [0, %d)", *dev, DLI_MAX_DEVICES);
  return DLI_OK;
}

int sm_count() {
  static int n[DLI_MAX_DEVICES] = {};
  int dev = 0;
  if (current_device(&dev) != DLI_OK) return 148;
  if (n[dev] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    n[dev] = v;
  }
  return n[dev];
}

bool pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = exp_env("DLI_NO_PDL");
    v = (e && e[0] == '1') ? 0 : 1;
  }
  return v == 1;
}

static thread_local bool g_barrier_next = false;
void pdl_barrier_next_launch() { g_barrier_next = true; }
bool pdl_take_barrier() {
  const bool b = g_barrier_next;
  g_barrier_next = false;
  return b;
}

int kv_view(const dli_kv_desc* kv, int layer, KvView* out) {
  DLI_REQUIRE(kv && kv->k_pages && kv->v_pages && kv->block_table && kv->seq_lens, "kv descriptor has a null pointer");
  DLI_REQUIRE(kv->page_tokens > 0 && kv->n_pages > 0 && kv->max_pages > 0 && kv->max_ctx > 0,
              "kv descriptor: page_tokens=%d n_pages=%d max_pages=%d max_ctx=%d", kv->page_tokens, kv->n_pages,
              kv->max_pages, kv->max_ctx);
  DLI_REQUIRE(layer >= 0, "negative layer index");
  out->k = reinterpret_cast<__nv_bfloat16*>(static_cast<char*>(kv->k_pages) + (long long)layer * kv->layer_stride_bytes);
  out->v = reinterpret_cast<__nv_bfloat16*>(static_cast<char*>(kv->v_pages) + (long long)layer * kv->layer_stride_bytes);
  out->block_table = kv->block_table;
  out->seq_lens = kv->seq_lens;
  out->page_tokens = kv->page_tokens;
  out->n_pages = kv->n_pages;
  out->max_pages = kv->max_pages;
  out->max_ctx = kv->max_ctx;
  return DLI_OK;
}

// ---- span object ------------------------------------------------------------------------------------
struct WeightMaps {
  CUtensorMap box64, box128, box256;
};
struct W8Maps {
  CUtensorMap box64, box128;  // uint8 maps of a quantised matrix
  const void* q = nullptr;
  const float* scale = nullptr;
  long long rows = 0, cols = 0;
};
struct Layer {
  dli_layer_weights w{};
  bool bound = false;
  bool w8 = false;            // weights bound through dli_span_bind_weights_w8 (row-wise int8 + fp32 scales)
  WeightMaps qkv, o, gu, down;
  W8Maps qkv8, o8, gu8, down8;
};
struct ScratchMaps {          // bf16 maps over the dequantisation scratch for one matrix shape (prefill of a w8 span)
  long long rows, cols;
  WeightMaps m;
};
struct ProfMark {
  cudaEvent_t ev;
  int cls;  // DLI_PROF_* of the launches that follow this mark; -1 closes a forward
};
struct ActMapKey {
  const void* ptr;
  int rows, cols, box;
  CUtensorMap map;
};

}  // namespace dli

struct dli_span {
  dli_span_desc d;
  int n_qkv;  // (H + 2 KV) * hd
  std::vector<dli::Layer> layers;
  std::vector<dli::ActMapKey> act_maps;  // activation-side TMA descriptors, keyed by (pointer, shape, box)
  void* dq_scratch = nullptr;            // caller-allocated bf16 scratch a w8 span dequantises into at prefill
  size_t dq_scratch_bytes = 0;
  std::vector<dli::ScratchMaps> dq_maps;
  dli_act_i8 a8{};                       // int8 activations (dli_span_set_act_i8): the one quantised operand of the GEMM about to run
  float a8_threshold = 0.f;
  bool a8_ready = false;                 // the RMSNorm launch just issued has already left its output quantised in a8
  std::vector<dli::ActMapKey> a8_maps;   // uint8 maps over a8.q, keyed by (rows, cols, box)
  const void* head_ptr = nullptr;        // lm_head matrix the maps below describe (dli_stage_decode_step)
  int head_vocab = 0;
  dli::WeightMaps head;
  // optional per-kernel-class timing (dli_span_profile_*): one event at every class boundary of the forward
  bool profiling = false;
  std::vector<dli::ProfMark> marks;
  size_t marks_used = 0;
};

namespace dli {

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

// Layout: the split-K accumulation buffer comes first, at an offset that does not depend on the call's
// shape, because it carries an invariant across calls (all zeros on entry and on exit); everything behind
// them is per-call scratch.
struct Workspace {
  float* qkv_acc;  // [128, n_qkv] fp32, zero between calls
  float* resid;
  __nv_bfloat16* xn;
  float* qkv;
  __nv_bfloat16* qbuf;
  __nv_bfloat16* attn;
  __nv_bfloat16* act;
  float* dec_scratch;
  size_t dec_scratch_bytes;
  size_t total;
};

static Workspace carve(const dli_span* s, void* base, int max_tokens, int max_batch, int max_ctx) {
  const dli_span_desc& d = s->d;
  char* p = static_cast<char*>(base);
  size_t off = 0;
  auto take = [&](size_t bytes) {
    char* r = p ? p + off : nullptr;
    off += align256(bytes);
    return r;
  };
  Workspace w{};
  const size_t M = (size_t)max_tokens;
  w.qkv_acc = (float*)take((size_t)128 * s->n_qkv * 4);
  w.resid = (float*)take(M * d.hidden_size * 4);
  w.xn = (__nv_bfloat16*)take(M * d.hidden_size * 2);
  w.qkv = (float*)take(max_tokens > 128 ? M * s->n_qkv * 4 : 0);
  w.qbuf = (__nv_bfloat16*)take(M * d.n_heads * d.head_dim * 2);
  w.attn = (__nv_bfloat16*)take(M * d.n_heads * d.head_dim * 2);
  w.act = (__nv_bfloat16*)take(M * d.intermediate_size * 2);
  w.dec_scratch_bytes = decode_attn_scratch_bytes(max_batch, d.n_heads, d.head_dim, max_ctx);
  w.dec_scratch = (float*)take(w.dec_scratch_bytes);
  w.total = off;
  return w;
}

static int act_map(dli_span* s, const void* ptr, int rows, int cols, int box, const CUtensorMap** out) {
  for (auto& e : s->act_maps)
    if (e.ptr == ptr && e.rows == rows && e.cols == cols && e.box == box) {
      *out = &e.map;
      return DLI_OK;
    }
  if (s->act_maps.size() >= 64) s->act_maps.clear();
  ActMapKey k{ptr, rows, cols, box, {}};
  int rc = make_tmap_bf16(&k.map, ptr, rows, cols, cols, box);
  if (rc) return rc;
  s->act_maps.push_back(k);
  *out = &s->act_maps.back().map;
  return DLI_OK;
}

static int prof_mark(dli_span* s, int cls, cudaStream_t st) {
  if (!s->profiling) return DLI_OK;
  if (s->marks_used == s->marks.size()) {
    ProfMark m{nullptr, -1};
    DLI_CUDA_OK(cudaEventCreate(&m.ev));
    s->marks.push_back(m);
  }
  ProfMark& m = s->marks[s->marks_used++];
  m.cls = cls;
  DLI_CUDA_OK(cudaEventRecord(m.ev, st));
  return DLI_OK;
}
#define DLI_MARK(cls) do { if ((rc = prof_mark(s, cls, st))) return rc; } while (0)

// one linear layer of the span: out (+)= x[m, k] . w[n, k]^T through the planned tcgen05 kernel
static int span_gemm(dli_span* s, const void* x, const WeightMaps& wm, void* out, long long ldo, int m, int n, int k,
                     int epilogue, cudaStream_t st, const RopeArgs* rope = nullptr, const W8Maps* w8 = nullptr);

// a quantised matrix at prefill: widen it into the scratch, then the bf16 tensor-core GEMM on the copy
static int span_gemm_dequant(dli_span* s, const void* x, const W8Maps& w8, void* out, long long ldo, int m, int n, int k,
                             int epilogue, cudaStream_t st, const RopeArgs* rope) {
  const size_t need = (size_t)w8.rows * w8.cols * 2;
  if (!s->dq_scratch || s->dq_scratch_bytes < need)
    return fail(DLI_ERR_WORKSPACE, "int8-weight span: chunks of more than 128 tokens need a dequantisation scratch of %zu bytes "
                                   "(dli_span_set_dequant_scratch)", need);
  int rc = launch_dequantize_rows_w8(w8.q, w8.scale, w8.rows, w8.cols, s->dq_scratch, st);
  if (rc) return rc;
  ScratchMaps* sm = nullptr;
  for (auto& e : s->dq_maps)
    if (e.rows == w8.rows && e.cols == w8.cols) sm = &e;
  if (!sm) {
    ScratchMaps e{w8.rows, w8.cols, {}};
    if ((rc = make_tmap_bf16(&e.m.box128, s->dq_scratch, w8.rows, w8.cols, w8.cols, 128))) return rc;
    if ((rc = make_tmap_bf16(&e.m.box64, s->dq_scratch, w8.rows, w8.cols, w8.cols, 64))) return rc;
    if ((rc = make_tmap_bf16(&e.m.box256, s->dq_scratch, w8.rows, w8.cols, w8.cols, 256))) return rc;
    s->dq_maps.push_back(e);
    sm = &s->dq_maps.back();
  }
  // the GEMM kernels request their weight tiles BEFORE the dependency wait (weights never change) -- not true of a
  // scratch the previous launch has just filled: this one GEMM starts only when the dequantisation has completed
  pdl_barrier_next_launch();
  return span_gemm(s, x, sm->m, out, ldo, m, n, k, epilogue, st, rope, nullptr);
}

static int span_gemm(dli_span* s, const void* x, const WeightMaps& wm, void* out, long long ldo, int m, int n, int k,
                     int epilogue, cudaStream_t st, const RopeArgs* rope, const W8Maps* w8) {
  if (w8 && w8->q) {
    if (m > 128 || epilogue == DLI_EPI_QKV_ROPE_INTERNAL) return span_gemm_dequant(s, x, *w8, out, ldo, m, n, k, epilogue, st, rope);
    if (s->a8.q) {  // int8 activations: quantise the operand (one CTA per token), then the kind::i8 GEMM
      static const bool skip_quant = exp_env("DLI_A8_SKIP_QUANT") != nullptr;   // experiments build only (results are garbage)
      int rc8 = (s->a8_ready || skip_quant) ? DLI_OK : launch_quantize_act_i8(x, k, m, k, s->a8_threshold, &s->a8, st);
      s->a8_ready = false;
      if (rc8) return rc8;
      const GemmPlan gi = plan_gemm_i8(m, n, k, epilogue);
      auto xq_map = [&](const void* base, const CUtensorMap** out_map) -> int {
        for (auto& e : s->a8_maps)
          if (e.ptr == base && e.rows == m && e.cols == k && e.box == gi.b_box_rows) { *out_map = &e.map; return DLI_OK; }
        ActMapKey e{base, m, k, gi.b_box_rows, {}};
        int r = make_tmap_u8(&e.map, base, m, k, gi.b_box_rows);
        if (r) return r;
        s->a8_maps.push_back(e);
        *out_map = &s->a8_maps.back().map;
        return DLI_OK;
      };
      const CUtensorMap *t1 = nullptr, *t2 = nullptr;
      if ((rc8 = xq_map(s->a8.q, &t1))) return rc8;
      t2 = t1;
      if (s->a8.q2) {
        if ((rc8 = xq_map(s->a8.q2, &t2))) return rc8;
        if ((rc8 = xq_map(s->a8.q, &t1))) return rc8;   // the vector may have grown: take the address again
      }
      return launch_gemm_i8_planned(gi, *t1, *t2, gi.a_box_rows == 64 ? w8->box64 : w8->box128, s->a8, w8->scale, out, ldo, m, n,
                                    epilogue, st);
    }
    const GemmPlan g8 = plan_gemm(m, n, k, epilogue);
    const CUtensorMap* tx8 = nullptr;
    int rc8 = act_map(s, x, m, k, g8.b_box_rows, &tx8);
    if (rc8) return rc8;
    return launch_gemm_w8_planned(g8, *tx8, g8.a_box_rows == 64 ? w8->box64 : w8->box128, w8->scale, out, ldo, m, n, epilogue, st);
  }
  const GemmPlan g = plan_gemm(m, n, k, epilogue);
  const int x_box = g.swap ? g.b_box_rows : g.a_box_rows;
  const int w_box = g.swap ? g.a_box_rows : g.b_box_rows;
  const CUtensorMap* tx = nullptr;
  int rc = act_map(s, x, m, k, x_box, &tx);
  if (rc) return rc;
  const CUtensorMap& tw = w_box == 256 ? wm.box256 : w_box == 64 ? wm.box64 : wm.box128;
  return launch_gemm_planned(g, *tx, tw, out, ldo, m, n, epilogue, st, rope);
}

}  // namespace dli

using namespace dli;

extern "C" {

int dli_version(void) { return DLI_VERSION; }
const char* dli_last_error(void) { return g_err; }
int dli_last_launch_count(void) { return g_launches; }

int dli_rmsnorm(const void* x, int x_is_f32, const void* weight_bf16, void* y_bf16, float* resid_out, int rows, int h,
                float eps, float pre_scale, dli_stream_t stream) {
  return launch_rmsnorm(x, x_is_f32 != 0, (const __nv_bfloat16*)weight_bf16, (__nv_bfloat16*)y_bf16, resid_out, nullptr,
                        rows, h, eps, pre_scale, (cudaStream_t)stream);
}

int dli_gemm(const void* x_bf16, int64_t ldx, const void* w_bf16, void* out, int64_t ldo, int m, int n, int k,
             int epilogue, dli_stream_t stream) {
  return launch_gemm(x_bf16, ldx, w_bf16, out, ldo, m, n, k, epilogue, (cudaStream_t)stream);
}

int dli_rope_kv_append(float* qkv_f32, int zero_after, const float* cos, const float* sin, void* q_bf16,
                       const dli_kv_desc* kv, int layer, int batch, int q_len, int n_heads, int n_kv_heads,
                       int head_dim, dli_stream_t stream) {
  KvView v;
  int rc = kv_view(kv, layer, &v);
  if (rc) return rc;
  return launch_rope_kv_append(qkv_f32, zero_after != 0, cos, sin, (__nv_bfloat16*)q_bf16, v, batch, q_len, n_heads,
                               n_kv_heads, head_dim, (cudaStream_t)stream);
}

size_t dli_decode_attention_scratch_bytes(int batch, int n_heads, int head_dim, int max_ctx) {
  return decode_attn_scratch_bytes(batch, n_heads, head_dim, max_ctx);
}

int dli_decode_attention(const void* q_bf16, void* out_bf16, const dli_kv_desc* kv, int layer, int batch, int n_heads,
                         int n_kv_heads, int head_dim, float* scratch, size_t scratch_bytes, dli_stream_t stream) {
  KvView v;
  int rc = kv_view(kv, layer, &v);
  if (rc) return rc;
  return launch_decode_attention((const __nv_bfloat16*)q_bf16, (__nv_bfloat16*)out_bf16, v, batch, n_heads, n_kv_heads,
                                 head_dim, scratch, scratch_bytes, (cudaStream_t)stream);
}

int dli_prefill_attention(const void* q_bf16, void* out_bf16, const dli_kv_desc* kv, int layer, int batch, int q_len,
                          int n_heads, int n_kv_heads, int head_dim, int causal, dli_stream_t stream) {
  KvView v;
  int rc = kv_view(kv, layer, &v);
  if (rc) return rc;
  return launch_prefill_attention((const __nv_bfloat16*)q_bf16, (__nv_bfloat16*)out_bf16, v, batch, q_len, n_heads,
                                  n_kv_heads, head_dim, causal != 0, (cudaStream_t)stream);
}

int dli_kv_evict(const dli_kv_desc* kv, int n_layers, int batch, int n_kv_heads, int head_dim, int cur_len, int window,
                 int sinks, int n_new, const float* cos_hist, const float* sin_hist, dli_stream_t stream) {
  return launch_kv_evict(kv, n_layers, batch, n_kv_heads, head_dim, cur_len, window, sinks, n_new, cos_hist, sin_hist,
                         (cudaStream_t)stream);
}

int dli_silu_mul(float* gu_f32, void* act_bf16, int m, int n, int zero_after, dli_stream_t stream) {
  return launch_silu_mul(gu_f32, (__nv_bfloat16*)act_bf16, m, n, zero_after != 0, (cudaStream_t)stream);
}

int dli_cast_f32_bf16(const float* in, void* out_bf16, int64_t n, dli_stream_t stream) {
  return launch_cast(in, (__nv_bfloat16*)out_bf16, n, (cudaStream_t)stream);
}

int dli_span_create(const dli_span_desc* desc, dli_span** out) {
  DLI_REQUIRE(desc && out, "dli_span_create: null pointer");
  DLI_REQUIRE(desc->dtype == DLI_BF16, "dli_span_create: only DLI_BF16 is implemented");
  DLI_REQUIRE(desc->hidden_size > 0 && desc->hidden_size % 8 == 0 && desc->intermediate_size > 0 &&
                  desc->intermediate_size % 8 == 0,
              "dli_span_create: hidden/intermediate sizes must be positive multiples of 8");
  DLI_REQUIRE(desc->n_heads > 0 && desc->n_kv_heads > 0 && desc->n_heads % desc->n_kv_heads == 0,
              "dli_span_create: n_heads must be a multiple of n_kv_heads");
  DLI_REQUIRE(desc->head_dim == 64 || desc->head_dim == 128, "dli_span_create: head_dim must be 64 or 128");
  const int G = desc->n_heads / desc->n_kv_heads;
  DLI_REQUIRE(G == 1 || G == 2 || G == 4 || G == 8, "dli_span_create: GQA group size %d not supported (1, 2, 4, 8)", G);
  DLI_REQUIRE(desc->n_layers > 0, "dli_span_create: n_layers must be positive");
  dli_span* s = new (std::nothrow) dli_span();
  if (!s) return fail(DLI_ERR_ARG, "dli_span_create: out of host memory");
  s->d = *desc;
  s->n_qkv = (desc->n_heads + 2 * desc->n_kv_heads) * desc->head_dim;
  s->layers.resize(desc->n_layers);
  s->act_maps.reserve(64);
  *out = s;
  return DLI_OK;
}

int dli_span_profile(dli_span* s, int enable) {
  DLI_REQUIRE(s, "dli_span_profile: null span");
  s->profiling = enable != 0;
  s->marks_used = 0;
  return DLI_OK;
}

int dli_span_profile_read(dli_span* s, float* ms_by_class, int* intervals_by_class) {
  DLI_REQUIRE(s && ms_by_class && intervals_by_class, "dli_span_profile_read: null pointer");
  for (int c = 0; c < DLI_PROF_CLASSES; ++c) { ms_by_class[c] = 0.f; intervals_by_class[c] = 0; }
  if (s->marks_used == 0) return DLI_OK;
  DLI_CUDA_OK(cudaEventSynchronize(s->marks[s->marks_used - 1].ev));
  for (size_t i = 0; i + 1 < s->marks_used; ++i) {
    const int c = s->marks[i].cls;
    if (c < 0 || c >= DLI_PROF_CLASSES) continue;
    float ms = 0.f;
    DLI_CUDA_OK(cudaEventElapsedTime(&ms, s->marks[i].ev, s->marks[i + 1].ev));
    ms_by_class[c] += ms;
    intervals_by_class[c] += 1;
  }
  s->marks_used = 0;
  return DLI_OK;
}

int dli_span_destroy(dli_span* span) {
  if (span) for (auto& m : span->marks) cudaEventDestroy(m.ev);
  delete span;
  return DLI_OK;
}

int dli_span_bind_weights(dli_span* s, int layer, const dli_layer_weights* w) {
  DLI_REQUIRE(s && w, "dli_span_bind_weights: null pointer");
  DLI_REQUIRE(layer >= 0 && layer < s->d.n_layers, "dli_span_bind_weights: layer %d out of range", layer);
  DLI_REQUIRE(w->wqkv && w->wo && w->wgu && w->wdown && w->g_in && w->g_post, "dli_span_bind_weights: null weight pointer");
  Layer& L = s->layers[layer];
  const dli_span_desc& d = s->d;
  const int qdim = d.n_heads * d.head_dim;
  struct { const void* p; long long rows, cols; WeightMaps* m; } items[4] = {
      {w->wqkv, s->n_qkv, d.hidden_size, &L.qkv},
      {w->wo, d.hidden_size, qdim, &L.o},
      {w->wgu, 2LL * d.intermediate_size, d.hidden_size, &L.gu},
      {w->wdown, d.hidden_size, d.intermediate_size, &L.down}};
  for (auto& it : items) {
    int rc = make_tmap_bf16(&it.m->box128, it.p, it.rows, it.cols, it.cols, 128);
    if (rc) return rc;
    rc = make_tmap_bf16(&it.m->box64, it.p, it.rows, it.cols, it.cols, 64);
    if (rc) return rc;
    rc = make_tmap_bf16(&it.m->box256, it.p, it.rows, it.cols, it.cols, 256);
    if (rc) return rc;
  }
  L.w = *w;
  L.bound = true;
  return DLI_OK;
}

int dli_span_bind_weights_w8(dli_span* s, int layer, const dli_layer_weights_w8* w) {
  DLI_REQUIRE(s && w, "dli_span_bind_weights_w8: null pointer");
  DLI_REQUIRE(layer >= 0 && layer < s->d.n_layers, "dli_span_bind_weights_w8: layer %d out of range", layer);
  DLI_REQUIRE(w->wqkv && w->wo && w->wgu && w->wdown && w->sqkv && w->so && w->sgu && w->sdown && w->g_in && w->g_post,
              "dli_span_bind_weights_w8: null weight or scale pointer");
  Layer& L = s->layers[layer];
  const dli_span_desc& d = s->d;
  const int qdim = d.n_heads * d.head_dim;
  struct { const void* p; const float* sc; long long rows, cols; W8Maps* m; } items[4] = {
      {w->wqkv, w->sqkv, s->n_qkv, d.hidden_size, &L.qkv8},
      {w->wo, w->so, d.hidden_size, qdim, &L.o8},
      {w->wgu, w->sgu, 2LL * d.intermediate_size, d.hidden_size, &L.gu8},
      {w->wdown, w->sdown, d.hidden_size, d.intermediate_size, &L.down8}};
  for (auto& it : items) {
    int rc = make_tmap_u8(&it.m->box128, it.p, it.rows, it.cols, 128);
    if (rc) return rc;
    if ((rc = make_tmap_u8(&it.m->box64, it.p, it.rows, it.cols, 64))) return rc;
    it.m->q = it.p;
    it.m->scale = it.sc;
    it.m->rows = it.rows;
    it.m->cols = it.cols;
  }
  L.w = dli_layer_weights{nullptr, nullptr, nullptr, nullptr, w->g_in, w->g_post};
  L.w8 = true;
  L.bound = true;
  return DLI_OK;
}

int dli_span_set_dequant_scratch(dli_span* s, void* scratch_bf16, size_t bytes) {
  DLI_REQUIRE(s, "dli_span_set_dequant_scratch: null span");
  DLI_REQUIRE(!scratch_bf16 || (reinterpret_cast<uintptr_t>(scratch_bf16) & 255) == 0, "dli_span_set_dequant_scratch: 256-byte alignment");
  s->dq_scratch = scratch_bf16;
  s->dq_scratch_bytes = scratch_bf16 ? bytes : 0;
  s->dq_maps.clear();
  return DLI_OK;
}

size_t dli_span_dequant_scratch_bytes(const dli_span* s) {
  if (!s) return 0;
  const dli_span_desc& d = s->d;
  const size_t a = (size_t)s->n_qkv * d.hidden_size, b = (size_t)d.hidden_size * d.n_heads * d.head_dim,
               c = (size_t)2 * d.intermediate_size * d.hidden_size;
  return 2 * std::max(a, std::max(b, c));
}

static int a8_kmax(const dli_span* s) {
  const dli_span_desc& d = s->d;
  return std::max(d.hidden_size, std::max(d.n_heads * d.head_dim, d.intermediate_size));
}
size_t dli_span_act_i8_scratch_bytes(const dli_span* s) { return s ? act_i8_bytes(128, a8_kmax(s), 2) : 0; }

int dli_span_set_act_i8(dli_span* s, void* scratch, size_t bytes, float threshold) {
  DLI_REQUIRE(s, "dli_span_set_act_i8: null span");
  s->a8_maps.clear();
  s->a8 = dli_act_i8{};
  s->a8_threshold = threshold;
  if (!scratch) return DLI_OK;
  return act_i8_carve(scratch, bytes, 128, a8_kmax(s), threshold > 0.f ? 2 : 1, &s->a8);
}

size_t dli_act_i8_bytes(int m, int k, int planes) { return (m > 0 && k > 0 && (planes == 1 || planes == 2)) ? act_i8_bytes(m, k, planes) : 0; }
int dli_act_i8_carve(void* base, size_t bytes, int m, int k, int planes, dli_act_i8* out) {
  return act_i8_carve(base, bytes, m, k, planes, out);
}
int dli_quantize_act_i8(const void* x_bf16, int64_t ldx, int m, int k, float threshold, const dli_act_i8* dst, dli_stream_t stream) {
  return launch_quantize_act_i8(x_bf16, ldx, m, k, threshold, dst, (cudaStream_t)stream);
}
int dli_gemm_i8(const dli_act_i8* x, const void* q_u8, const float* w_scale, void* out, int64_t ldo, int m, int n, int k, int epilogue,
                dli_stream_t stream) {
  return launch_gemm_i8(x, q_u8, w_scale, out, ldo, m, n, k, epilogue, (cudaStream_t)stream);
}

int dli_gemm_w8(const void* x_bf16, int64_t ldx, const void* q_u8, const float* scale, void* out, int64_t ldo, int m, int n,
                int k, int epilogue, dli_stream_t stream) {
  return launch_gemm_w8(x_bf16, ldx, q_u8, scale, out, ldo, m, n, k, epilogue, (cudaStream_t)stream);
}

int dli_quantize_rows_w8(const void* w_bf16, int64_t n, int64_t k, void* q_u8, float* scale, dli_stream_t stream) {
  return launch_quantize_rows_w8(w_bf16, n, k, q_u8, scale, (cudaStream_t)stream);
}

int dli_dequantize_rows_w8(const void* q_u8, const float* scale, int64_t n, int64_t k, void* w_bf16, dli_stream_t stream) {
  return launch_dequantize_rows_w8(q_u8, scale, n, k, w_bf16, (cudaStream_t)stream);
}

size_t dli_span_workspace_bytes(const dli_span* span, int max_tokens, int max_batch, int max_ctx) {
  if (!span || max_tokens <= 0 || max_batch <= 0 || max_ctx <= 0) return 0;
  return carve(span, nullptr, max_tokens, max_batch, max_ctx).total;
}

static int span_forward_impl(dli_span* s, const dli_kv_desc* kv, int batch, int q_len, int causal, const void* hidden_in,
                             void* hidden_out, const float* cos, const float* sin, void* all_hidden, void* workspace,
                             size_t workspace_bytes, dli_stream_t stream, const StepState& step,
                             const dli_stage_io* io = nullptr, const dli_attn_mask* mask = nullptr);

int dli_span_forward(dli_span* s, const dli_kv_desc* kv, int batch, int q_len, int causal, const void* hidden_in,
                     void* hidden_out, const float* cos, const float* sin, void* all_hidden, void* workspace,
                     size_t workspace_bytes, dli_stream_t stream) {
  const StepState none{nullptr, nullptr, nullptr, 0, false};
  return span_forward_impl(s, kv, batch, q_len, causal, hidden_in, hidden_out, cos, sin, all_hidden, workspace,
                           workspace_bytes, stream, none);
}

int dli_span_forward_masked(dli_span* s, const dli_kv_desc* kv, int batch, int q_len, int causal, const dli_attn_mask* mask,
                            const void* hidden_in, void* hidden_out, const float* cos, const float* sin, void* all_hidden,
                            void* workspace, size_t workspace_bytes, dli_stream_t stream) {
  DLI_REQUIRE(mask && (mask->key_limit || mask->key_valid), "dli_span_forward_masked: empty mask (use dli_span_forward)");
  const StepState none{nullptr, nullptr, nullptr, 0, false};
  return span_forward_impl(s, kv, batch, q_len, causal, hidden_in, hidden_out, cos, sin, all_hidden, workspace,
                           workspace_bytes, stream, none, nullptr, mask);
}

int dli_prefill_attention_masked(const void* q_bf16, void* out_bf16, const dli_kv_desc* kv, int layer, int batch, int q_len,
                                 int n_heads, int n_kv_heads, int head_dim, int causal, const dli_attn_mask* mask,
                                 dli_stream_t stream) {
  DLI_REQUIRE(mask, "dli_prefill_attention_masked: null mask");
  KvView v;
  int rc = kv_view(kv, layer, &v);
  if (rc) return rc;
  return launch_prefill_attention_masked((const __nv_bfloat16*)q_bf16, (__nv_bfloat16*)out_bf16, v, batch, q_len, n_heads,
                                         n_kv_heads, head_dim, causal != 0, AttnMask{mask->key_limit, mask->key_valid, mask->key_stride},
                                         (cudaStream_t)stream);
}

int dli_decode_attention_masked(const void* q_bf16, void* out_bf16, const dli_kv_desc* kv, int layer, int batch, int n_heads,
                                int n_kv_heads, int head_dim, const dli_attn_mask* mask, float* scratch, size_t scratch_bytes,
                                dli_stream_t stream) {
  DLI_REQUIRE(mask, "dli_decode_attention_masked: null mask");
  KvView v;
  int rc = kv_view(kv, layer, &v);
  if (rc) return rc;
  return launch_decode_attention_masked((const __nv_bfloat16*)q_bf16, (__nv_bfloat16*)out_bf16, v, batch, n_heads, n_kv_heads,
                                        head_dim, AttnMask{nullptr, mask->key_valid, mask->key_stride}, scratch, scratch_bytes,
                                        (cudaStream_t)stream);
}

int dli_span_decode_step(dli_span* s, const dli_kv_desc* kv, int batch, const void* hidden_in, void* hidden_out,
                         const float* cos_table, const float* sin_table, long long* position, int advance,
                         void* workspace, size_t workspace_bytes, dli_stream_t stream) {
  DLI_REQUIRE(kv && position, "dli_span_decode_step: null pointer");
  const StepState step{position, static_cast<int*>(const_cast<void*>(static_cast<const void*>(kv->seq_lens))), position, batch,
                       advance != 0};
  return span_forward_impl(s, kv, batch, 1, 1, hidden_in, hidden_out, cos_table, sin_table, nullptr, workspace,
                           workspace_bytes, stream, step);
}

int dli_stage_decode_step(dli_span* s, const dli_kv_desc* kv, int batch, const dli_stage_io* io, const float* cos_table,
                          const float* sin_table, long long* position, int advance, void* workspace,
                          size_t workspace_bytes, dli_stream_t stream) {
  DLI_REQUIRE(s && kv && io && position, "dli_stage_decode_step: null pointer");
  DLI_REQUIRE(io->hidden_in || (io->token_ids && io->embed_table), "dli_stage_decode_step: no input (hidden_in, or token_ids + embed_table)");
  DLI_REQUIRE(io->hidden_out || io->lm_head, "dli_stage_decode_step: no output (hidden_out, or lm_head + tokens_out)");
  if (io->lm_head) {
    DLI_REQUIRE(io->final_norm && io->logits && io->tokens_out && io->vocab > 0 && io->vocab % 8 == 0,
                "dli_stage_decode_step: the head needs final_norm, logits, tokens_out and a vocabulary that is a multiple of 8");
    if (s->head_ptr != io->lm_head || s->head_vocab != io->vocab) {
      int rc;
      if ((rc = make_tmap_bf16(&s->head.box128, io->lm_head, io->vocab, s->d.hidden_size, s->d.hidden_size, 128))) return rc;
      if ((rc = make_tmap_bf16(&s->head.box64, io->lm_head, io->vocab, s->d.hidden_size, s->d.hidden_size, 64))) return rc;
      if ((rc = make_tmap_bf16(&s->head.box256, io->lm_head, io->vocab, s->d.hidden_size, s->d.hidden_size, 256))) return rc;
      s->head_ptr = io->lm_head;
      s->head_vocab = io->vocab;
    }
  }
  DLI_REQUIRE(!(io->wait_flag || io->signal_flag) || io->state, "dli_stage_decode_step: flags need the 4-word step state");
  const StepState step{position, static_cast<int*>(const_cast<void*>(static_cast<const void*>(kv->seq_lens))), position, batch,
                       advance != 0};
  return span_forward_impl(s, kv, batch, 1, 1, io->hidden_in, io->hidden_out, cos_table, sin_table, nullptr, workspace,
                           workspace_bytes, stream, step, io);
}

int dli_flag_wait(const uint32_t* flag, uint32_t at_least, uint32_t* state, int timeout_ms, dli_stream_t stream) {
  return launch_flag_wait(flag, at_least, state, (unsigned long long)(timeout_ms > 0 ? timeout_ms : 2000) * 1000000ull,
                          (cudaStream_t)stream);
}

int dli_lm_head_argmax(const void* hidden, int hidden_is_f32, const void* final_norm_bf16, float eps, const void* lm_head_bf16,
                       int vocab, int h, int batch, void* xn_bf16, float* logits, int32_t* tokens_out, dli_stream_t stream) {
  DLI_REQUIRE(hidden && final_norm_bf16 && lm_head_bf16 && xn_bf16 && logits && tokens_out, "dli_lm_head_argmax: null pointer");
  DLI_REQUIRE(batch > 0 && batch <= 128, "dli_lm_head_argmax: batch %d outside [1, 128]", batch);
  cudaStream_t st = (cudaStream_t)stream;
  int rc = launch_rmsnorm(hidden, hidden_is_f32 != 0, (const __nv_bfloat16*)final_norm_bf16, (__nv_bfloat16*)xn_bf16, nullptr,
                          nullptr, batch, h, eps, 1.0f, st);
  if (rc) return rc;
  if ((rc = launch_gemm(xn_bf16, h, lm_head_bf16, logits, vocab, batch, vocab, h, DLI_EPI_STORE_F32, st))) return rc;
  return launch_argmax(logits, tokens_out, batch, vocab, st, nullptr, ExitArgs{nullptr, nullptr});
}

static int span_forward_impl(dli_span* s, const dli_kv_desc* kv, int batch, int q_len, int causal, const void* hidden_in,
                             void* hidden_out, const float* cos, const float* sin, void* all_hidden, void* workspace,
                             size_t workspace_bytes, dli_stream_t stream, const StepState& step, const dli_stage_io* io,
                             const dli_attn_mask* mask) {
  reset_launch_count();
  DLI_REQUIRE(s && kv && cos && sin && workspace && (io || (hidden_in && hidden_out)), "dli_span_forward: null pointer");
  DLI_REQUIRE(batch > 0 && q_len > 0, "dli_span_forward: batch=%d q_len=%d", batch, q_len);
  DLI_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "dli_span_forward: workspace must be 256-byte aligned");
  const dli_span_desc& d = s->d;
  const int M = batch * q_len;
  const Workspace w = carve(s, workspace, M, batch, kv->max_ctx);
  if (w.total > workspace_bytes)
    return fail(DLI_ERR_WORKSPACE, "dli_span_forward: workspace %zu bytes < %zu needed", workspace_bytes, w.total);
  for (int j = 0; j < d.n_layers; ++j)
    DLI_REQUIRE(s->layers[j].bound, "dli_span_forward: layer %d has no weights bound", j);
  cudaStream_t st = (cudaStream_t)stream;
  const int h = d.hidden_size, inter = d.intermediate_size, qdim = d.n_heads * d.head_dim;
  const bool skinny = M <= 128;  // weight-streaming GEMM path accumulates into pre-zeroed fp32 buffers
  // experiment builds only (DLI_EXPERIMENTS; results are garbage): bit c set in DLI_DEBUG_SKIP drops every launch of
  // dli_prof_class c, which measures what that class costs inside the PDL-chained, graph-replayed step
#ifdef DLI_EXPERIMENTS
  static const unsigned skip_mask = getenv("DLI_DEBUG_SKIP") ? (unsigned)strtoul(getenv("DLI_DEBUG_SKIP"), nullptr, 0) : 0u;
#define DLI_SKIPPED(cls) (skip_mask & (1u << (cls)))
#else
#define DLI_SKIPPED(cls) false
#endif
  __nv_bfloat16* hs = static_cast<__nv_bfloat16*>(all_hidden);
  int rc;

  // one launch class of the step: timing mark, optional skip (DLI_DEBUG_SKIP), error propagation
#define DLI_STEP(cls, call)                  \
  do {                                       \
    DLI_MARK(cls);                           \
    rc = DLI_SKIPPED(cls) ? DLI_OK : (call); \
    if (rc) return rc;                       \
  } while (0)
  const __nv_bfloat16* attn_in = w.qbuf;

  // the call's first kernel starts after everything already in the stream -- unless the caller of a stage step vouches
  // that what precedes it in the stream belongs to OTHER generations (micro-batches of a pipeline round): then the
  // steps are chained like the kernels inside one
  if (!(io && io->chain)) pdl_barrier_next_launch();
  // int8 activations: the two RMSNorm launches of a layer leave their output quantised for the GEMM that follows
  const QuantArgs qnorm{s->a8, s->a8_threshold, (h + 127) / 128};
  s->a8_ready = false;
  for (int j = 0; j < d.n_layers; ++j) {
    Layer& L = s->layers[j];
    KvView view;
    if ((rc = kv_view(kv, j, &view))) return rc;
    const QuantArgs* qn = (L.w8 && s->a8.q && skinny && rmsnorm_can_quantize(h)) ? &qnorm : nullptr;
    // input_layernorm (modules.py:162); layer 0 also opens the fp32 residual stream (and is never skipped)
    DLI_MARK(DLI_PROF_RMSNORM);
    if (j == 0 && io) {  // pipeline stage step: flag wait and/or embedding gather inside the entry kernel
      const EntryArgs ea{io->token_ids, io->wait_flag, io->state, io->wait_lag,
                         (unsigned long long)(io->timeout_ms > 0 ? io->timeout_ms : 2000) * 1000000ull};
      rc = launch_rmsnorm_entry(io->token_ids ? io->embed_table : hidden_in, (const __nv_bfloat16*)L.w.g_in, w.xn, w.resid, M, h,
                                d.eps_input, ea, st, qn);
      if (rc) return rc;
      s->a8_ready = qn != nullptr;
    } else if (j == 0 || !DLI_SKIPPED(DLI_PROF_RMSNORM)) {
      rc = launch_rmsnorm(j == 0 ? hidden_in : (const void*)w.resid, j != 0, (const __nv_bfloat16*)L.w.g_in, w.xn,
                          j == 0 ? w.resid : nullptr, hs ? hs + (size_t)j * M * h : nullptr, M, h, d.eps_input, 1.0f, st, qn);
      if (rc) return rc;
      s->a8_ready = qn != nullptr;
    }
    // q|k|v projection (modules.py:61-63), RoPE and cache append (modules.py:73-85, cache.py:103-109):
    //   prefill  RoPE + append are the epilogue of the projection GEMM (whole heads per 256-column tile);
    //   decode   split-K GEMM into the fp32 accumulation buffer, then either the RoPE/append kernel (which re-zeroes
    //            the buffer behind itself) or, when the context is not split, the attention kernel does that work
    //            itself (rope_inline.cuh) -- one launch less per layer.
    const bool fused_rope = !skinny && s->n_qkv % 256 == 0 && gemm_persistent_enabled();
    const bool fused_dec_rope = !fused_rope && q_len == 1 && skinny && !mask &&
                                decode_attention_fuses_rope(view, batch, d.n_heads, d.n_kv_heads, d.head_dim);
    if (fused_rope) {
      RopeArgs ra{cos, sin, w.qbuf, view, q_len, d.n_heads, d.n_kv_heads, d.head_dim};
      DLI_STEP(DLI_PROF_GEMM_QKV, span_gemm(s, w.xn, L.qkv, nullptr, 0, M, s->n_qkv, h, DLI_EPI_QKV_ROPE_INTERNAL, st, &ra, L.w8 ? &L.qkv8 : nullptr));
    } else {
      float* qkv = skinny ? w.qkv_acc : w.qkv;
      DLI_STEP(DLI_PROF_GEMM_QKV, span_gemm(s, w.xn, L.qkv, qkv, s->n_qkv, M, s->n_qkv, h,
                                            skinny ? DLI_EPI_ADD_F32 : DLI_EPI_STORE_F32, st, nullptr, L.w8 ? &L.qkv8 : nullptr));
      DLI_MARK(DLI_PROF_ROPE_APPEND);
      if (!fused_dec_rope && !DLI_SKIPPED(DLI_PROF_ROPE_APPEND)) {
        rc = launch_rope_kv_append(qkv, skinny, cos, sin, w.qbuf, view, batch, q_len, d.n_heads, d.n_kv_heads, d.head_dim,
                                   st, step.pos);
        if (rc) return rc;
      }
    }
    // attention core (modules.py:87-105)
    if (mask) {  // padded keys / per-query key limits (model.py:103-135 beyond the causal flag): the generic kernels
      const AttnMask am{mask->key_limit, mask->key_valid, mask->key_stride};
      if (q_len == 1)
        DLI_STEP(DLI_PROF_ATTENTION, launch_decode_attention_masked(attn_in, w.attn, view, batch, d.n_heads, d.n_kv_heads,
                                                                    d.head_dim, am, w.dec_scratch, w.dec_scratch_bytes, st));
      else
        DLI_STEP(DLI_PROF_ATTENTION, launch_prefill_attention_masked(attn_in, w.attn, view, batch, q_len, d.n_heads,
                                                                     d.n_kv_heads, d.head_dim, causal != 0, am, st));
    } else if (q_len == 1) {
      const DecRope dr{w.qkv_acc, cos, sin, s->n_qkv, step.pos};
      DLI_STEP(DLI_PROF_ATTENTION,
               launch_decode_attention(attn_in, w.attn, view, batch, d.n_heads, d.n_kv_heads, d.head_dim, w.dec_scratch,
                                       w.dec_scratch_bytes, st, /*early_kv=*/true, fused_dec_rope ? &dr : nullptr));
    } else {
      DLI_STEP(DLI_PROF_ATTENTION, launch_prefill_attention(attn_in, w.attn, view, batch, q_len, d.n_heads, d.n_kv_heads,
                                                            d.head_dim, causal != 0, st));
    }
    // o_proj + residual (modules.py:112,173)
    DLI_STEP(DLI_PROF_GEMM_O, span_gemm(s, w.attn, L.o, w.resid, h, M, h, qdim, DLI_EPI_ADD_F32, st, nullptr, L.w8 ? &L.o8 : nullptr));
    // post_attention_layernorm of h + h (modules.py:179)
    DLI_STEP(DLI_PROF_RMSNORM, launch_rmsnorm(w.resid, true, (const __nv_bfloat16*)L.w.g_post, w.xn, nullptr, nullptr, M, h,
                                              d.eps_post, d.post_prescale, st, qn));
    s->a8_ready = qn != nullptr && !DLI_SKIPPED(DLI_PROF_RMSNORM);
    // MLP (modules.py:181,182): gate|up with silu(gate) * up as the GEMM epilogue on both paths (64 gate + 64 up weight
    // rows per CTA at decode, 128 + 128 accumulator columns per CTA pair at prefill), then down + residual
    DLI_STEP(DLI_PROF_GEMM_GATE_UP, span_gemm(s, w.xn, L.gu, w.act, inter, M, inter, h, DLI_EPI_SILU_MUL_BF16, st, nullptr, L.w8 ? &L.gu8 : nullptr));
    DLI_STEP(DLI_PROF_GEMM_DOWN, span_gemm(s, w.act, L.down, w.resid, h, M, h, inter, DLI_EPI_ADD_F32, st, nullptr, L.w8 ? &L.down8 : nullptr));
  }
#undef DLI_STEP
  DLI_MARK(DLI_PROF_CAST);
  if (io) {  // pipeline stage step: the LAST kernel advances the device-resident state and signals the consumer
    const ExitArgs xa{io->signal_flag, io->state}, none{nullptr, nullptr};
    const bool head = io->lm_head != nullptr;
    if (hidden_out && (rc = launch_cast(w.resid, static_cast<__nv_bfloat16*>(hidden_out), (long long)M * h, st,
                                        head ? nullptr : &step, head ? &none : &xa)))
      return rc;
    if (head) {  // last stage of a token ring: final RMSNorm -> lm_head -> greedy argmax (the client's tail, on the GPU)
      if ((rc = launch_rmsnorm(w.resid, true, (const __nv_bfloat16*)io->final_norm, w.xn, nullptr, nullptr, M, h, io->eps_final,
                               1.0f, st)))
        return rc;
      if ((rc = span_gemm(s, w.xn, s->head, io->logits, io->vocab, M, io->vocab, h, DLI_EPI_STORE_F32, st))) return rc;
      if ((rc = launch_argmax(io->logits, io->tokens_out, M, io->vocab, st, &step, xa))) return rc;
    }
    DLI_MARK(-1);
    return DLI_OK;
  }
  if ((rc = launch_cast(w.resid, static_cast<__nv_bfloat16*>(hidden_out), (long long)M * h, st, hs ? nullptr : &step)))
    return rc;
  if (hs && (rc = launch_cast(w.resid, hs + (size_t)d.n_layers * M * h, (long long)M * h, st, &step))) return rc;
  DLI_MARK(-1);
  return DLI_OK;
}

}  // extern "C"

#ifdef DLI_TRACE
// trace builds only (DLI_TRACE=1 build.sh): point every translation unit's timeline buffer at `buf`
// (word 0 = event counter, word 1 = capacity in events, then 2 words per event); null disables
namespace dli {
int trace_set_gemm(void*);
int trace_set_norm(void*);
int trace_set_attn(void*);
int trace_set_attn_gqa(void*);
}  // namespace dli
extern "C" int dli_debug_trace(void* buf) {
  return dli::trace_set_gemm(buf) | dli::trace_set_norm(buf) | dli::trace_set_attn(buf) | dli::trace_set_attn_gqa(buf);
}
#endif
