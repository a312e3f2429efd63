// Shared host/device helpers for libdli_b200 (sm_100a only).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "../../include/dli_b200.h"

namespace dli {

// ---- host-side error plumbing (thread-local message, never throws across the ABI) -----------------
void set_error(const char* fmt, ...);
int fail(int code, const char* fmt, ...);
void count_launch(int n = 1);
void reset_launch_count();

#define DLI_CUDA_OK(expr)                                                                      \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess)                                                                     \
      return ::dli::fail(DLI_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                         __FILE__, __LINE__);                                                  \
  } while (0)

#define DLI_CHECK_LAUNCH(name)                                                                 \
  do {                                                                                         \
    cudaError_t _e = cudaGetLastError();                                                       \
    if (_e != cudaSuccess)                                                                     \
      return ::dli::fail(DLI_ERR_CUDA, "launch of %s failed: %s", name, cudaGetErrorString(_e)); \
    ::dli::count_launch();                                                                     \
  } while (0)

#define DLI_REQUIRE(cond, ...)                                   \
  do {                                                           \
    if (!(cond)) return ::dli::fail(DLI_ERR_ARG, __VA_ARGS__);   \
  } while (0)

int sm_count();  // SMs of the current device (cached per device)
bool pdl_enabled();  // programmatic dependent launch for the span's kernel chain

// A/B and timing switches read from the environment exist only in experiment builds (DLI_EXPERIMENTS=1 build.sh ->
// libdli_b200_exp.so, loaded with DLI_LIB); the product library never looks at the environment.
#ifdef DLI_EXPERIMENTS
inline const char* exp_env(const char* name) { return getenv(name); }
#else
inline const char* exp_env(const char*) { return nullptr; }
#endif

// One-time set-up that CUDA keeps PER DEVICE (function attributes such as the > 48 KB dynamic shared memory
// opt-in): `slot` is a static array owned by the call site (one per kernel instantiation).
constexpr int DLI_MAX_DEVICES = 64;
struct OncePerDevice {
  bool done[DLI_MAX_DEVICES] = {};
};
int current_device(int* dev);  // cudaGetDevice with range check against DLI_MAX_DEVICES
#define DLI_ONCE_PER_DEVICE(expr)                        \
  do {                                                   \
    static ::dli::OncePerDevice _once;                   \
    int _dev = 0;                                        \
    if (int _rc = ::dli::current_device(&_dev)) return _rc; \
    if (!_once.done[_dev]) {                             \
      DLI_CUDA_OK(expr);                                 \
      _once.done[_dev] = true;                           \
    }                                                    \
  } while (0)
// The next launch_kernel() of this thread is launched WITHOUT the attribute, i.e. it starts only when everything
// before it in the stream has completed.  dli_span_forward does this for its first kernel, which bounds how far
// back a pre-wait prologue can overlap: nothing launched before the current forward call is still running when
// any of the call's kernels executes, so prologues may read what earlier calls wrote (the KV pages of past tokens).
void pdl_barrier_next_launch();
bool pdl_take_barrier();

// Launch with the programmatic-stream-serialization attribute: the kernel may begin (prologue, weight prefetch)
// while its predecessor in the stream is still running; it must execute pdl_wait() before touching anything the
// predecessor produces or still reads.
template <typename... KArgs, typename... Args>
int launch_kernel(const char* name, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                  Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = (pdl_enabled() && !pdl_take_barrier()) ? 1 : 0;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
  if (e != cudaSuccess) return fail(DLI_ERR_CUDA, "launch of %s failed: %s", name, cudaGetErrorString(e));
  count_launch();
  return DLI_OK;
}

// ---- internal launchers shared between the op-level ABI and the span runner ------------------------
struct KvView {
  __nv_bfloat16* k;  // this layer's K pool [n_pages][KV][page_tokens][hd]
  __nv_bfloat16* v;
  const int32_t* block_table;
  const int32_t* seq_lens;
  int page_tokens, n_pages, max_pages, max_ctx;
};
int kv_view(const dli_kv_desc* kv, int layer, KvView* out);

// Arguments of the fused q|k|v projection epilogue (prefill): RoPE + paged KV append straight from the accumulator.
struct RopeArgs {
  const float* cos;  // [q_len, hd]
  const float* sin;
  __nv_bfloat16* q_out;  // [tokens, H * hd]
  KvView kv;
  int q_len, H, KV, hd;
};
constexpr int DLI_EPI_QKV_ROPE_INTERNAL = 4;  // not part of the public enum: only the span runner uses it
bool gemm_persistent_enabled();

// ---- GEMM planning / launch (gemm_tcgen05.cu) --------------------------------------------------------
struct GemmPlan {
  bool swap;          // weights are the 128-row MMA operand (decode); else tokens are
  int bn;             // MMA N
  int a_box_rows, b_box_rows;  // TMA box rows of the A-role / B-role operand
  int k_blocks, kb_per_split, splits, stages;
  int grid_x, grid_y;
  bool pair;          // prefill only: CTA-pair kernel (cta_group::2), B box = 128 rows per CTA
};
GemmPlan plan_gemm(int m, int n, int k, int epilogue);
int make_tmap_bf16(CUtensorMap* map, const void* base, long long rows, long long cols, long long ld, int box_rows);
int launch_gemm_planned(const GemmPlan& g, const CUtensorMap& tmap_x, const CUtensorMap& tmap_w, void* out,
                        long long ldo, int m, int n, int epilogue, cudaStream_t st, const RopeArgs* rope = nullptr);
int launch_gemm(const void* x, long long ldx, const void* w, void* out, long long ldo, int m, int n, int k,
                int epilogue, cudaStream_t st);
// weight-only int8 (gemm_w8.cu): decode GEMM on row-wise quantised weights, (de)quantisation kernels
int make_tmap_u8(CUtensorMap* map, const void* base, long long rows, long long cols, int box_rows);
int launch_gemm_w8_planned(const GemmPlan& g, const CUtensorMap& tmap_x, const CUtensorMap& tmap_w, const float* scale, void* out,
                           long long ldo, int m, int n, int epilogue, cudaStream_t st);
int launch_gemm_w8(const void* x, long long ldx, const void* q, const float* scale, void* out, long long ldo, int m, int n,
                   int k, int epilogue, cudaStream_t st);
int launch_quantize_rows_w8(const void* w_bf16, long long n, long long k, void* q_u8, float* scale, cudaStream_t st);
int launch_dequantize_rows_w8(const void* q_u8, const float* scale, long long n, long long k, void* w_bf16, cudaStream_t st);
// int8 x int8 (gemm_i8.cu): activation quantiser and the kind::i8 decode GEMM
GemmPlan plan_gemm_i8(int m, int n, int k, int epilogue);
int launch_gemm_i8_planned(const GemmPlan& g, const CUtensorMap& tmap_x, const CUtensorMap& tmap_x2, const CUtensorMap& tmap_w,
                           const dli_act_i8& x, const float* w_scale, void* out, long long ldo, int m, int n, int epilogue,
                           cudaStream_t st);
int launch_gemm_i8(const dli_act_i8* x, const void* q, const float* w_scale, void* out, long long ldo, int m, int n, int k,
                   int epilogue, cudaStream_t st);
int launch_quantize_act_i8(const void* x_bf16, long long ldx, int m, int k, float threshold, const dli_act_i8* dst, cudaStream_t st);
size_t act_i8_bytes(int m, int k, int planes);
int act_i8_carve(void* base, size_t bytes, int m, int k, int planes, dli_act_i8* out);

// Decode step whose position lives on the device (dli_span_decode_step): which table row to use, and what the span's
// last kernel advances afterwards.
struct StepState {
  const long long* pos;  // device: position of the step's token; nullptr = cos/sin already point at its row
  int* seq_lens;         // device [batch]: += 1 by the last kernel when advance
  long long* pos_rw;     // device: += 1 by the last kernel when advance
  int batch;
  bool advance;
};

// Hand-off of a pipeline stage step, folded into the step's own first and last kernels (dli_stage_decode_step):
//   entry  the first kernel spins on a flag in THIS GPU's memory until the producer (previous stage, or the last stage
//          through the ring closure) has stored this step's input; with token_ids the input rows are gathered from an
//          embedding table (stage 0 of a token ring);
//   exit   the last kernel, after every CTA's stores (possibly into a peer GPU's memory) are fenced system-wide,
//          bumps the step counter and releases the consumer's flag.
// state: 4 words in this GPU's memory: [0] steps completed, [1] CTA ticket (0 between steps), [2] sticky error
// (a wait gave up after timeout_ns), [3] unused.
struct EntryArgs {
  const int* token_ids;       // nullable
  const uint32_t* wait_flag;  // nullable
  uint32_t* state;            // nullable when wait_flag is
  int lag;                    // wait until *wait_flag + lag >= state[0] + 1
  unsigned long long timeout_ns;
};
struct ExitArgs {
  uint32_t* signal_flag;  // nullable
  uint32_t* state;        // nullable: no counter, no signal
};

// Attention masks beyond the causal flag (dli_attn_mask in include/dli_b200.h), passed by value to the kernels.
struct AttnMask {
  const int32_t* key_limit;   // device [q_len] or nullptr: query l sees cache slots j < key_limit[l]
  const uint8_t* key_valid;   // device [batch, key_stride] or nullptr: 0 = padded key, visible to no query
  int key_stride;
};
int launch_prefill_attention_masked(const __nv_bfloat16* q, __nv_bfloat16* out, const KvView& kv, int batch, int q_len, int H,
                                    int KV, int hd, bool causal, const AttnMask& mask, cudaStream_t st);
int launch_decode_attention_masked(const __nv_bfloat16* q, __nv_bfloat16* out, const KvView& kv, int batch, int H, int KV,
                                   int hd, const AttnMask& mask, float* scratch, size_t scratch_bytes, cudaStream_t st);

// ---- other launchers ----------------------------------------------------------------------------------
// RMSNorm launches that also leave their output quantised for the int8 x int8 GEMM that follows (norm_rope.cu)
struct QuantArgs {
  dli_act_i8 d;      // d.q == nullptr: no quantised output
  float threshold;
  int k_blocks;      // ceil(h / 128)
};
bool rmsnorm_can_quantize(int h);
int launch_rmsnorm_entry(const void* x, const __nv_bfloat16* w, __nv_bfloat16* y, float* resid_out, int rows, int h, float eps,
                         const EntryArgs& ea, cudaStream_t st, const QuantArgs* quant = nullptr);
int launch_argmax(const float* logits, int32_t* tokens_out, int batch, int vocab, cudaStream_t st, const StepState* adv,
                  const ExitArgs& xa);
int launch_flag_wait(const uint32_t* flag, uint32_t at_least, uint32_t* state, unsigned long long timeout_ns, cudaStream_t st);
int launch_rmsnorm(const void* x, bool x_is_f32, const __nv_bfloat16* w, __nv_bfloat16* y, float* resid_out,
                   __nv_bfloat16* copy_out, int rows, int h, float eps, float pre_scale, cudaStream_t st,
                   const QuantArgs* quant = nullptr);
int launch_rope_kv_append(float* qkv, bool zero_after, const float* cos, const float* sin, __nv_bfloat16* q_out,
                          const KvView& kv, int batch, int q_len, int H, int KV, int hd, cudaStream_t st,
                          const long long* pos = nullptr);  // pos: cos/sin are tables, first row = *pos
int launch_silu_mul(float* gu, __nv_bfloat16* act, int m, int n, bool zero_after, cudaStream_t st);
int launch_kv_evict(const dli_kv_desc* kv, int n_layers, int batch, int KV, int hd, int cur_len, int window, int sinks,
                    int n_new, const float* cos_hist, const float* sin_hist, cudaStream_t st);
int launch_cast(const float* in, __nv_bfloat16* out, long long n, cudaStream_t st, const StepState* adv = nullptr,
                const ExitArgs* xa = nullptr);
size_t decode_attn_scratch_bytes(int batch, int H, int hd, int max_ctx);
// Decode attention that applies RoPE and the KV append itself (rope_inline.cuh); qkv == nullptr: q arrives rotated in
// bf16 and the pages already hold the step's token.  Only for unsplit contexts: decode_attention_fuses_rope().
struct DecRope {
  const float* qkv;  // [batch, n_all] fp32 q|k|v projection of the step (split-K accumulation buffer; re-zeroed)
  const float* cos;  // [hd] of the step's position
  const float* sin;
  int n_all;         // (H + 2 KV) * hd
  const long long* pos;  // nullable: cos/sin are whole tables and the step's row is *pos (device memory)
};
bool decode_attention_fuses_rope(const KvView& kv, int batch, int H, int KV, int hd);
bool decode_attn_gqa_supported(const KvView& kv, int G, int hd);
int launch_decode_attention_gqa(const __nv_bfloat16* q, __nv_bfloat16* out, const KvView& kv, int batch, int H, int KV,
                                int hd, int splits, float* scratch, bool early_kv, const DecRope& rope, cudaStream_t st);
// early_kv: the K/V pages of every token but the step's own were written before the current chain of launches began
// (see pdl_barrier_next_launch), so the kernel may stream them before its dependency wait.  The span runner says yes;
// the op-level ABI, which knows nothing about its caller's stream, says no.
int launch_decode_attention(const __nv_bfloat16* q, __nv_bfloat16* out, const KvView& kv, int batch, int H, int KV,
                            int hd, float* scratch, size_t scratch_bytes, cudaStream_t st, bool early_kv = false,
                            const DecRope* rope = nullptr);
bool prefill_attn_tc_supported(const KvView& kv, int hd);
bool prefill_attn_tc2_supported(const KvView& kv, int hd, int q_len);
int launch_prefill_attention_tc2(const __nv_bfloat16* q, __nv_bfloat16* out, const KvView& kv, int batch, int q_len, int H,
                                 int KV, bool causal, cudaStream_t st);
int launch_prefill_attention_tc(const __nv_bfloat16* q, __nv_bfloat16* out, const KvView& kv, int batch, int q_len, int H,
                                int KV, bool causal, cudaStream_t st);
int launch_prefill_attention(const __nv_bfloat16* q, __nv_bfloat16* out, const KvView& kv, int batch, int q_len,
                             int H, int KV, int hd, bool causal, cudaStream_t st);

// ---- device helpers -------------------------------------------------------------------------------
#ifdef __CUDACC__
// Programmatic dependent launch: wait until the previous kernel of the stream has completed and its writes are
// visible / let the next kernel of the stream start its prologue.  Both are no-ops without the launch attribute.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
// ---- in-kernel timeline (only in builds made with DLI_TRACE=1 build.sh; see tests/exp_trace.py) ------------
// Every CTA stamps %globaltimer at kernel entry (phase 0), when griddepcontrol.wait returns (1) and at exit (2)
// into a host-provided buffer: word 0 = event counter, word 1 = capacity, then (tag, time) pairs.
#ifdef DLI_TRACE
static __device__ unsigned long long* g_trace_buf = nullptr;
__device__ __forceinline__ void trace_ev(int kid, int phase) {
  unsigned long long* t = g_trace_buf;
  if (!t) return;
  unsigned long long now;
  unsigned smid;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
  asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
  const unsigned long long i = atomicAdd(t, 1ull);
  if (i >= t[1]) return;
  const unsigned cta = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
  t[2 + 2 * i] = ((unsigned long long)kid << 56) | ((unsigned long long)phase << 52) | ((unsigned long long)smid << 32) | cta;
  t[3 + 2 * i] = now;
}
#define DLI_TRACE_EV(kid, phase) trace_ev(kid, phase)
#define DLI_TRACE_SETTER(name)                                                       \
  int name(void* buf) {                                                              \
    unsigned long long* p = static_cast<unsigned long long*>(buf);                   \
    return cudaMemcpyToSymbol(g_trace_buf, &p, sizeof(p)) == cudaSuccess ? 0 : 1;    \
  }
#else
#define DLI_TRACE_EV(kid, phase) ((void)0)
#define DLI_TRACE_SETTER(name)
#endif
enum { TR_GEMM = 1, TR_RMSNORM = 2, TR_ROPE = 3, TR_ATTN = 4, TR_COMBINE = 5, TR_CAST = 6, TR_ATTN_GQA = 7 };

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float bf16_lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }

// LLM.int8 activation codes of 8 consecutive entries of a row (gemm_i8.cu): entries below `thr` go to the first plane
// (code = round(v * inv)), the others to the outlier plane (round(v * inv2)); an entry is 0 in the plane it does not
// belong to.  p1 / p2: the 8 codes of each plane as bytes; s1 / s2: their sums.
__device__ __forceinline__ void i8_codes8(const float (&v)[8], float thr, float inv, float inv2, uint2& p1, uint2& p2, int& s1, int& s2) {
  uint32_t o[2] = {0, 0}, o2[2] = {0, 0};
  s1 = 0;
  s2 = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const bool out = fabsf(v[j]) >= thr;
    const int c = max(-127, min(127, __float2int_rn(v[j] * (out ? inv2 : inv))));
    const int c1 = out ? 0 : c, c2 = out ? c : 0;
    s1 += c1;
    s2 += c2;
    o[j >> 2] |= ((uint32_t)c1 & 0xffu) << (8 * (j & 3));
    o2[j >> 2] |= ((uint32_t)c2 & 0xffu) << (8 * (j & 3));
  }
  p1 = make_uint2(o[0], o[1]);
  p2 = make_uint2(o2[0], o2[1]);
}
// ---- cross-GPU flags (system scope) ---------------------------------------------------------------------
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// 128-bit load of data another GPU may have stored while this kernel was already running (a hand-off inbox): never
// through the non-coherent path, never from a stale L1 line
__device__ __forceinline__ uint4 ld_relaxed_sys_v4(const void* p) {
  uint4 r;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// One thread: wait until *flag has reached `need` (counters only grow; compared modulo 2^32).  Gives up after
// timeout_ns -- or at once if an earlier wait on this state already gave up -- and records that in state[2], so a
// producer that never signals turns into an error the host can read, not a hung GPU.
__device__ __forceinline__ void flag_wait(const uint32_t* flag, uint32_t need, uint32_t* state, unsigned long long timeout_ns) {
  if ((int)(ld_acquire_sys(flag) - need) >= 0) return;
  const unsigned long long t0 = global_timer_ns();
  unsigned ns = 32;
  while (true) {
    if ((int)(ld_acquire_sys(flag) - need) >= 0) return;
    if (ld_relaxed_sys(state + 2) != 0) return;
    if (global_timer_ns() - t0 > timeout_ns) {
      atomicExch(state + 2, 1u);
      return;
    }
    __nanosleep(ns);
    if (ns < 128) ns *= 2;
  }
}
// Last kernel of a stage step, called by ALL threads of every CTA after their output stores: fence them system-wide,
// take a ticket, and let the last CTA to arrive bump the step counter and release the consumer's flag.
__device__ __forceinline__ void step_exit(const ExitArgs& xa) {
  if (!xa.state) return;
  __syncthreads();  // every thread's output stores are ordered before thread 0's fence (the pattern of a grid-wide sync)
  if (threadIdx.x == 0) {
    __threadfence_system();  // cumulative: covers what the CTA's other threads stored before the barrier
    const unsigned total = gridDim.x * gridDim.y * gridDim.z;
    const unsigned ticket = atomicAdd(xa.state + 1, 1u);
    if (ticket == total - 1) {
      __threadfence_system();
      xa.state[1] = 0;
      const uint32_t done = xa.state[0] + 1;
      xa.state[0] = done;
      if (xa.signal_flag) st_release_sys(xa.signal_flag, done);
    }
  }
}

// 128-bit streaming load that does not allocate in L1 (data read once: weights, KV pages)
__device__ __forceinline__ uint4 ld_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
#endif

}  // namespace dli
