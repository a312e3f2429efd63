// HBM-bound element-wise kernels of the span: RMSNorm, RoPE + paged KV append, SiLU*up, fp32->bf16 cast.
#include "common.cuh"

#include <algorithm>

namespace dli {

// ---------------------------------------------------------------------------------------------------
// RMSNorm (LlamaRMSNorm.forward as called at modules.py:162 and :179).  One CTA per row; fp32 math;
// y = bf16(w * (s x) * rsqrt(mean((s x)^2) + eps)).  s = 2 reproduces the reference's norm of h + h.
// Optional side outputs: fp32 copy (bf16 hand-off -> fp32 residual stream at span entry) and bf16 copy
// (output_hidden_states, model.py:57-61).
// ---------------------------------------------------------------------------------------------------
constexpr int NORM_THREADS = 256;

template <bool F32IN, bool COHERENT = false>
__device__ __forceinline__ void load8(const void* x, long long i, float (&v)[8]) {
  if (COHERENT) {  // bf16 rows another GPU may have stored while this kernel was already resident (hand-off inbox)
    const uint4 a = ld_relaxed_sys_v4((const __nv_bfloat16*)x + i);
    v[0] = bf16_lo(a.x); v[1] = bf16_hi(a.x); v[2] = bf16_lo(a.y); v[3] = bf16_hi(a.y);
    v[4] = bf16_lo(a.z); v[5] = bf16_hi(a.z); v[6] = bf16_lo(a.w); v[7] = bf16_hi(a.w);
  } else if (F32IN) {
    const float4 a = *reinterpret_cast<const float4*>((const float*)x + i);
    const float4 b = *reinterpret_cast<const float4*>((const float*)x + i + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  } else {
    const uint4 a = *reinterpret_cast<const uint4*>((const __nv_bfloat16*)x + i);
    v[0] = bf16_lo(a.x); v[1] = bf16_hi(a.x); v[2] = bf16_lo(a.y); v[3] = bf16_hi(a.y);
    v[4] = bf16_lo(a.z); v[5] = bf16_hi(a.z); v[6] = bf16_lo(a.w); v[7] = bf16_hi(a.w);
  }
}

// One CTA per row.  Rows of up to NORM_THREADS * 8 * NORM_CHUNKS (8192) elements are read ONCE and kept in
// registers; the norm weights do not depend on the previous kernel, so they are fetched before the dependency
// wait and the chain after it is a single load, one block reduction and the store.  (At decode this kernel is
// pure latency between two weight-streaming GEMMs: every microsecond here is one the next GEMM's prefetched
// ring cannot cover.)  Wider rows take the two-pass path.
constexpr int NORM_CHUNKS = 4;

// ENTRY (first kernel of a pipeline stage step, bf16 input): the input rows are either a hand-off inbox that the
// previous stage fills over NVLink -- thread 0 waits for the producer's flag after the dependency wait, the rows are
// then read with coherent loads -- or rows of an embedding table picked by device-resident token ids (stage 0 of a
// token ring; the ids are what the flag guards).
template <bool F32IN, bool ENTRY = false, bool QUANT = false>
__global__ void __launch_bounds__(NORM_THREADS) rmsnorm_kernel(const void* __restrict__ x, const __nv_bfloat16* __restrict__ w,
                                                               __nv_bfloat16* __restrict__ y, float* __restrict__ resid_out,
                                                               __nv_bfloat16* __restrict__ copy_out, int h, float eps,
                                                               float s, EntryArgs ea, QuantArgs qa) {
  __shared__ float red[NORM_THREADS / 32];
  __shared__ float red_q[2][NORM_THREADS / 32];
  pdl_trigger();
  if (threadIdx.x == 0) DLI_TRACE_EV(TR_RMSNORM, 0);
  long long in_base = (long long)blockIdx.x * h;
  const long long base = in_base;
  const int i0 = threadIdx.x * 8;
  if (h <= NORM_THREADS * 8 * NORM_CHUNKS) {
    uint4 wv[NORM_CHUNKS];
    float v[NORM_CHUNKS][8];
#pragma unroll
    for (int c = 0; c < NORM_CHUNKS; ++c) {
      const int i = i0 + c * NORM_THREADS * 8;
      if (i < h) wv[c] = *reinterpret_cast<const uint4*>(w + i);
    }
    pdl_wait();
    if (threadIdx.x == 0) DLI_TRACE_EV(TR_RMSNORM, 1);
    if (ENTRY) {
      if (ea.wait_flag) {
        if (threadIdx.x == 0) flag_wait(ea.wait_flag, ld_relaxed_sys(ea.state) + 1u - (uint32_t)ea.lag, ea.state, ea.timeout_ns);
        __syncthreads();
      }
      if (ea.token_ids) in_base = (long long)ld_relaxed_sys(reinterpret_cast<const uint32_t*>(ea.token_ids) + blockIdx.x) * h;
    }
    float ss = 0.f;
#pragma unroll
    for (int c = 0; c < NORM_CHUNKS; ++c) {
      const int i = i0 + c * NORM_THREADS * 8;
      if (i < h) {
        if (ENTRY && !ea.token_ids) load8<F32IN, true>(x, in_base + i, v[c]);
        else load8<F32IN>(x, in_base + i, v[c]);
      }
    }
#pragma unroll
    for (int c = 0; c < NORM_CHUNKS; ++c) {
      const int i = i0 + c * NORM_THREADS * 8;
      if (i < h) {
#pragma unroll
        for (int j = 0; j < 8; ++j) ss += v[c][j] * v[c][j];
        if (resid_out) {
          *reinterpret_cast<float4*>(resid_out + base + i) = make_float4(v[c][0], v[c][1], v[c][2], v[c][3]);
          *reinterpret_cast<float4*>(resid_out + base + i + 4) = make_float4(v[c][4], v[c][5], v[c][6], v[c][7]);
        }
        if (copy_out) {
          *reinterpret_cast<uint4*>(copy_out + base + i) = make_uint4(pack_bf16(v[c][0], v[c][1]), pack_bf16(v[c][2], v[c][3]),
                                                                      pack_bf16(v[c][4], v[c][5]), pack_bf16(v[c][6], v[c][7]));
        }
      }
    }
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < NORM_THREADS / 32; ++k) t += red[k];  // same order in every thread: one value per row
    const float r = rsqrtf(t * s * s / (float)h + eps) * s;
    uint32_t ob[NORM_CHUNKS][4];
#pragma unroll
    for (int c = 0; c < NORM_CHUNKS; ++c) {
      const int i = i0 + c * NORM_THREADS * 8;
      if (i < h) {
        const float g[8] = {bf16_lo(wv[c].x), bf16_hi(wv[c].x), bf16_lo(wv[c].y), bf16_hi(wv[c].y),
                            bf16_lo(wv[c].z), bf16_hi(wv[c].z), bf16_lo(wv[c].w), bf16_hi(wv[c].w)};
#pragma unroll
        for (int j = 0; j < 4; ++j) ob[c][j] = pack_bf16(g[2 * j] * (v[c][2 * j] * r), g[2 * j + 1] * (v[c][2 * j + 1] * r));
        *reinterpret_cast<uint4*>(y + base + i) = make_uint4(ob[c][0], ob[c][1], ob[c][2], ob[c][3]);
      }
    }
    if (QUANT) {   // its own instantiation: the 16-bit path runs the kernel without this tail
      // the row also leaves quantised for the int8 x int8 GEMM that follows (what quantize_act_i8_kernel would make of y)
      const float thr = (qa.d.q2 != nullptr && qa.threshold > 0.f) ? qa.threshold : INFINITY;
      float mx = 0.f, mo = 0.f;
#pragma unroll
      for (int c = 0; c < NORM_CHUNKS; ++c) {
        if (i0 + c * NORM_THREADS * 8 < h) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float v0 = fabsf(bf16_lo(ob[c][j])), v1 = fabsf(bf16_hi(ob[c][j]));
            if (v0 < thr) mx = fmaxf(mx, v0); else mo = fmaxf(mo, v0);
            if (v1 < thr) mx = fmaxf(mx, v1); else mo = fmaxf(mo, v1);
          }
        }
      }
      mx = warp_max(mx);
      mo = warp_max(mo);
      if ((threadIdx.x & 31) == 0) { red_q[0][threadIdx.x >> 5] = mx; red_q[1][threadIdx.x >> 5] = mo; }
      __syncthreads();
      mx = red_q[0][0];
      mo = red_q[1][0];
#pragma unroll
      for (int k = 1; k < NORM_THREADS / 32; ++k) { mx = fmaxf(mx, red_q[0][k]); mo = fmaxf(mo, red_q[1][k]); }
      const float inv = mx > 0.f ? 127.0f / mx : 0.f, inv2 = mo > 0.f ? 127.0f / mo : 0.f;
      const long long row = blockIdx.x;
      if (threadIdx.x == 0) {
        qa.d.scale[row] = mx / 127.0f;
        if (qa.d.q2) qa.d.scale2[row] = mo / 127.0f;
      }
      int8_t* dst = (int8_t*)qa.d.q + row * h;
      int8_t* dst2 = (int8_t*)qa.d.q2 + row * h;
      int32_t* ks = qa.d.ksum + row * qa.k_blocks;
      int32_t* ks2 = qa.d.ksum2 + row * qa.k_blocks;
#pragma unroll
      for (int c = 0; c < NORM_CHUNKS; ++c) {   // 16 consecutive threads share a 128-entry block; every lane takes part in the shuffles
        const int i = i0 + c * NORM_THREADS * 8;
        int sum = 0, sum2 = 0;
        if (i < h) {
          const float f[8] = {bf16_lo(ob[c][0]), bf16_hi(ob[c][0]), bf16_lo(ob[c][1]), bf16_hi(ob[c][1]),
                              bf16_lo(ob[c][2]), bf16_hi(ob[c][2]), bf16_lo(ob[c][3]), bf16_hi(ob[c][3])};
          uint2 p1, p2;
          i8_codes8(f, thr, inv, inv2, p1, p2, sum, sum2);
          *reinterpret_cast<uint2*>(dst + i) = p1;
          if (qa.d.q2) *reinterpret_cast<uint2*>(dst2 + i) = p2;
        }
#pragma unroll
        for (int off = 8; off > 0; off >>= 1) {
          sum += __shfl_xor_sync(0xffffffffu, sum, off);
          sum2 += __shfl_xor_sync(0xffffffffu, sum2, off);
        }
        if ((threadIdx.x & 15) == 0 && i < h) {
          ks[i >> 7] = sum;
          if (qa.d.q2) ks2[i >> 7] = sum2;
        }
      }
    }
    if (threadIdx.x == 0) DLI_TRACE_EV(TR_RMSNORM, 2);
    return;
  }
  // ---- wide rows: two passes (the second one hits L1/L2)
  __shared__ float rinv;
  pdl_wait();
  float ss = 0.f;
  float v[8];
  for (int i = i0; i < h; i += NORM_THREADS * 8) {
    load8<F32IN>(x, base + i, v);
#pragma unroll
    for (int j = 0; j < 8; ++j) ss += v[j] * v[j];
    if (resid_out) {
      *reinterpret_cast<float4*>(resid_out + base + i) = make_float4(v[0], v[1], v[2], v[3]);
      *reinterpret_cast<float4*>(resid_out + base + i + 4) = make_float4(v[4], v[5], v[6], v[7]);
    }
    if (copy_out) {
      *reinterpret_cast<uint4*>(copy_out + base + i) =
          make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
    }
  }
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < NORM_THREADS / 32 ? red[threadIdx.x] : 0.f;
    t = warp_sum(t);
    if (threadIdx.x == 0) rinv = rsqrtf(t * s * s / (float)h + eps);
  }
  __syncthreads();
  const float r = rinv * s;
  for (int i = i0; i < h; i += NORM_THREADS * 8) {
    load8<F32IN>(x, base + i, v);
    const uint4 wv = *reinterpret_cast<const uint4*>(w + i);
    const float g[8] = {bf16_lo(wv.x), bf16_hi(wv.x), bf16_lo(wv.y), bf16_hi(wv.y),
                        bf16_lo(wv.z), bf16_hi(wv.z), bf16_lo(wv.w), bf16_hi(wv.w)};
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) o[j] = pack_bf16(g[2 * j] * (v[2 * j] * r), g[2 * j + 1] * (v[2 * j + 1] * r));
    *reinterpret_cast<uint4*>(y + base + i) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// quant: also leave the rows quantised for the int8 x int8 GEMM (rows of up to 8192 entries only: rmsnorm_can_quantize)
bool rmsnorm_can_quantize(int h) { return h <= NORM_THREADS * 8 * NORM_CHUNKS && h % 16 == 0; }

int launch_rmsnorm(const void* x, bool x_is_f32, const __nv_bfloat16* w, __nv_bfloat16* y, float* resid_out,
                   __nv_bfloat16* copy_out, int rows, int h, float eps, float pre_scale, cudaStream_t st, const QuantArgs* quant) {
  DLI_REQUIRE(x && w && y, "dli_rmsnorm: null pointer");
  DLI_REQUIRE(rows > 0 && h > 0 && h % 8 == 0, "dli_rmsnorm: rows=%d h=%d (h must be a positive multiple of 8)", rows, h);
  DLI_REQUIRE(!quant || rmsnorm_can_quantize(h), "rmsnorm with quantised output: h = %d not supported", h);
  auto kern = quant ? (x_is_f32 ? rmsnorm_kernel<true, false, true> : rmsnorm_kernel<false, false, true>)
                    : (x_is_f32 ? rmsnorm_kernel<true> : rmsnorm_kernel<false>);
  return launch_kernel("rmsnorm_kernel", kern, dim3(rows), dim3(NORM_THREADS), 0, st, x, w, y, resid_out, copy_out, h, eps, pre_scale,
                       EntryArgs{}, quant ? *quant : QuantArgs{});
}

int launch_rmsnorm_entry(const void* x, const __nv_bfloat16* w, __nv_bfloat16* y, float* resid_out, int rows, int h, float eps,
                         const EntryArgs& ea, cudaStream_t st, const QuantArgs* quant) {
  DLI_REQUIRE(x && w && y && resid_out, "stage step entry: null pointer");
  DLI_REQUIRE(rows > 0 && h > 0 && h % 8 == 0 && h <= NORM_THREADS * 8 * NORM_CHUNKS,
              "stage step entry: rows=%d h=%d (h must be a multiple of 8, at most %d)", rows, h, NORM_THREADS * 8 * NORM_CHUNKS);
  DLI_REQUIRE(!ea.wait_flag || ea.state, "stage step entry: a wait flag needs the step state");
  return launch_kernel("rmsnorm_kernel<entry>", quant ? rmsnorm_kernel<false, true, true> : rmsnorm_kernel<false, true>, dim3(rows),
                       dim3(NORM_THREADS), 0, st, x, w, y, resid_out, (__nv_bfloat16*)nullptr, h, eps, 1.0f, ea,
                       quant ? *quant : QuantArgs{});
}

// ---------------------------------------------------------------------------------------------------
// RoPE (modules.py:17-20: x*cos + rotate_half(x)*sin, rotate_half = cat(-x2, x1)) fused with the KV
// append (cache.py:103-109).  One CTA per token; each thread rotates two adjacent (d, d + hd/2) pairs.
// q goes to a dense bf16 [tokens, H*hd] buffer, rotated k and raw v go to the sequence's pages.
// ---------------------------------------------------------------------------------------------------
constexpr int ROPE_THREADS = 256;

__global__ void __launch_bounds__(ROPE_THREADS) rope_kv_append_kernel(float* __restrict__ qkv, int zero_after,
                                                                      const float* __restrict__ cosT,
                                                                      const float* __restrict__ sinT,
                                                                      __nv_bfloat16* __restrict__ q_out, KvView kv,
                                                                      int q_len, int H, int KV, int hd,
                                                                      const long long* __restrict__ pos0) {
  pdl_trigger();
  if (threadIdx.x == 0) DLI_TRACE_EV(TR_ROPE, 0);
  const int t = blockIdx.x;
  const int b = t / q_len, l = t - b * q_len;
  const int half = hd >> 1;
  const int n_all = (H + 2 * KV) * hd;
  float* row = qkv + (long long)t * n_all;
  // sequence lengths and the block table are host-side state of the cache, never written by a kernel of the
  // chain this launch belongs to: look the destination slot up while the projection GEMM is still running
  const int pos = kv.seq_lens[b] + l;
  const int page = kv.block_table[(long long)b * kv.max_pages + pos / kv.page_tokens];
  const int slot = pos % kv.page_tokens;
  pdl_wait();
  if (threadIdx.x == 0) DLI_TRACE_EV(TR_ROPE, 1);
  const long long row0 = pos0 ? *pos0 : 0;  // tables indexed by absolute position, or rows already gathered
  const float* cs = cosT + (row0 + l) * hd;
  const float* sn = sinT + (row0 + l) * hd;

  const int pairs2 = (H + KV) * half / 2;  // work items: two adjacent d per thread
  const int tid = blockIdx.y * ROPE_THREADS + threadIdx.x, stride = gridDim.y * ROPE_THREADS;  // a token's work is cut over gridDim.y CTAs
  for (int it = tid; it < pairs2; it += stride) {
    const int p = it * 2;
    const int head = p / half, d = p - head * half;
    float* x = row + head * hd + d;
    const float2 x1 = *reinterpret_cast<const float2*>(x);
    const float2 x2 = *reinterpret_cast<const float2*>(x + half);
    const float2 c1 = *reinterpret_cast<const float2*>(cs + d), c2 = *reinterpret_cast<const float2*>(cs + d + half);
    const float2 s1 = *reinterpret_cast<const float2*>(sn + d), s2 = *reinterpret_cast<const float2*>(sn + d + half);
    const float o1x = x1.x * c1.x - x2.x * s1.x, o1y = x1.y * c1.y - x2.y * s1.y;
    const float o2x = x2.x * c2.x + x1.x * s2.x, o2y = x2.y * c2.y + x1.y * s2.y;
    __nv_bfloat16* dst;
    if (head < H) {
      dst = q_out + (long long)t * H * hd + head * hd + d;
    } else {
      dst = kv.k + (((long long)page * KV + (head - H)) * kv.page_tokens + slot) * hd + d;
    }
    *reinterpret_cast<uint32_t*>(dst) = pack_bf16(o1x, o1y);
    *reinterpret_cast<uint32_t*>(dst + half) = pack_bf16(o2x, o2y);
    if (zero_after) {
      *reinterpret_cast<float2*>(x) = make_float2(0.f, 0.f);
      *reinterpret_cast<float2*>(x + half) = make_float2(0.f, 0.f);
    }
  }
  float* vrow = row + (H + KV) * hd;
  for (int e = tid * 4; e < KV * hd; e += stride * 4) {
    const float4 a = *reinterpret_cast<const float4*>(vrow + e);
    const int kvh = e / hd, d = e - kvh * hd;
    __nv_bfloat16* dst = kv.v + (((long long)page * KV + kvh) * kv.page_tokens + slot) * hd + d;
    *reinterpret_cast<uint2*>(dst) = make_uint2(pack_bf16(a.x, a.y), pack_bf16(a.z, a.w));
    if (zero_after) *reinterpret_cast<float4*>(vrow + e) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  if (threadIdx.x == 0) DLI_TRACE_EV(TR_ROPE, 2);
}

int launch_rope_kv_append(float* qkv, bool zero_after, const float* cos, const float* sin, __nv_bfloat16* q_out,
                          const KvView& kv, int batch, int q_len, int H, int KV, int hd, cudaStream_t st,
                          const long long* pos) {
  DLI_REQUIRE(qkv && cos && sin && q_out, "dli_rope_kv_append: null pointer");
  DLI_REQUIRE(batch > 0 && q_len > 0, "dli_rope_kv_append: batch=%d q_len=%d", batch, q_len);
  DLI_REQUIRE(hd % 8 == 0 && H % KV == 0, "dli_rope_kv_append: head_dim %% 8 and H %% KV must be 0 (hd=%d H=%d KV=%d)", hd, H, KV);
  // few tokens (decode): spread each token over several CTAs so the kernel is not one latency chain per token
  const int items = (H + KV) * hd / 4;
  int ny = 1;
  if (batch * q_len < 2 * sm_count()) ny = std::max(1, std::min(8, items / ROPE_THREADS));
  return launch_kernel("rope_kv_append_kernel", rope_kv_append_kernel, dim3(batch * q_len, ny), dim3(ROPE_THREADS), 0, st, qkv,
                       zero_after ? 1 : 0, cos, sin, q_out, kv, q_len, H, KV, hd, pos);
}

// ---------------------------------------------------------------------------------------------------
// act = silu(gate) * up from the fp32 split-K accumulation buffer [m, 2n] (LlamaMLP.forward, modules.py:181)
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) silu_mul_kernel(float* __restrict__ gu, __nv_bfloat16* __restrict__ act, int m, int n,
                                                       int zero_after) {
  pdl_trigger();
  pdl_wait();
  const long long total4 = (long long)m * n / 4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += (long long)gridDim.x * blockDim.x) {
    const long long e = i * 4;
    const int r = (int)(e / n), c = (int)(e - (long long)r * n);
    float* g = gu + (long long)r * 2 * n + c;
    const float4 a = *reinterpret_cast<const float4*>(g);
    const float4 u = *reinterpret_cast<const float4*>(g + n);
    const float o0 = a.x / (1.f + __expf(-a.x)) * u.x, o1 = a.y / (1.f + __expf(-a.y)) * u.y;
    const float o2 = a.z / (1.f + __expf(-a.z)) * u.z, o3 = a.w / (1.f + __expf(-a.w)) * u.w;
    *reinterpret_cast<uint2*>(act + (long long)r * n + c) = make_uint2(pack_bf16(o0, o1), pack_bf16(o2, o3));
    if (zero_after) {
      *reinterpret_cast<float4*>(g) = make_float4(0.f, 0.f, 0.f, 0.f);
      *reinterpret_cast<float4*>(g + n) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
}

int launch_silu_mul(float* gu, __nv_bfloat16* act, int m, int n, bool zero_after, cudaStream_t st) {
  DLI_REQUIRE(gu && act, "dli_silu_mul: null pointer");
  DLI_REQUIRE(m > 0 && n > 0 && n % 4 == 0, "dli_silu_mul: m=%d n=%d (n must be a multiple of 4)", m, n);
  const long long total4 = (long long)m * n / 4;
  const int blocks = (int)std::min<long long>((total4 + 255) / 256, (long long)sm_count() * 8);
  return launch_kernel("silu_mul_kernel", silu_mul_kernel, dim3(blocks), dim3(256), 0, st, gu, act, m, n, zero_after ? 1 : 0);
}

__global__ void __launch_bounds__(256) cast_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, long long n4,
                                                   int* __restrict__ seq_lens, long long* __restrict__ pos, int batch,
                                                   ExitArgs xa) {
  pdl_trigger();
  pdl_wait();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 a = *reinterpret_cast<const float4*>(in + i * 4);
    *reinterpret_cast<uint2*>(out + i * 4) = make_uint2(pack_bf16(a.x, a.y), pack_bf16(a.z, a.w));
  }
  // Last kernel of a decode step whose bookkeeping lives on the device: every kernel that read the lengths or the
  // position is complete (the wait above is transitive), and the next step starts with a stream-ordered launch.
  if (seq_lens && blockIdx.x == 0) {
    for (int b = threadIdx.x; b < batch; b += blockDim.x) seq_lens[b] += 1;
    if (threadIdx.x == 0 && pos) *pos += 1;
  }
  step_exit(xa);  // pipeline stage: `out` may be the next stage's inbox on a peer GPU
}

int launch_cast(const float* in, __nv_bfloat16* out, long long n, cudaStream_t st, const StepState* adv, const ExitArgs* xa) {
  DLI_REQUIRE(in && out && n > 0 && n % 4 == 0, "dli_cast_f32_bf16: n=%lld must be a positive multiple of 4", n);
  const int blocks = (int)std::min<long long>((n / 4 + 255) / 256, (long long)sm_count() * 8);
  const bool a = adv && adv->advance;
  return launch_kernel("cast_kernel", cast_kernel, dim3(blocks), dim3(256), 0, st, in, out, n / 4,
                       a ? adv->seq_lens : (int*)nullptr, a ? adv->pos_rw : (long long*)nullptr, a ? adv->batch : 0,
                       xa ? *xa : ExitArgs{nullptr, nullptr});
}

// ---------------------------------------------------------------------------------------------------
// Greedy sampling at the end of the last pipeline stage: token[b] = argmax_v logits[b, v] (lowest index on ties).
// One CTA per sequence; the result may go straight into stage 0's token inbox on a peer GPU.  As the last kernel of
// that stage's step it also advances the device-resident lengths / position and signals the consumer.
// ---------------------------------------------------------------------------------------------------
constexpr int ARGMAX_THREADS = 512;

__global__ void __launch_bounds__(ARGMAX_THREADS) argmax_kernel(const float* __restrict__ logits, int32_t* __restrict__ tokens_out,
                                                                int vocab, int* __restrict__ seq_lens,
                                                                long long* __restrict__ pos, int batch, ExitArgs xa) {
  __shared__ float sv[ARGMAX_THREADS / 32];
  __shared__ int si[ARGMAX_THREADS / 32];
  pdl_trigger();
  pdl_wait();
  const float* row = logits + (long long)blockIdx.x * vocab;
  float best = -INFINITY;
  int arg = 0x7fffffff;
  const int v4 = vocab / 4;
  for (int i = threadIdx.x; i < v4; i += ARGMAX_THREADS) {
    const float4 a = *reinterpret_cast<const float4*>(row + 4 * i);
    const float e[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (e[j] > best) { best = e[j]; arg = 4 * i + j; }   // indices grow within a thread: first maximum wins
  }
  for (int i = 4 * v4 + threadIdx.x; i < vocab; i += ARGMAX_THREADS)
    if (row[i] > best) { best = row[i]; arg = i; }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
    if (ob > best || (ob == best && oa < arg)) { best = ob; arg = oa; }
  }
  if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = arg; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int k = 1; k < ARGMAX_THREADS / 32; ++k)
      if (sv[k] > best || (sv[k] == best && si[k] < arg)) { best = sv[k]; arg = si[k]; }
    tokens_out[blockIdx.x] = arg == 0x7fffffff ? 0 : arg;
  }
  if (seq_lens && blockIdx.x == 0) {
    for (int b = threadIdx.x; b < batch; b += blockDim.x) seq_lens[b] += 1;
    if (threadIdx.x == 0 && pos) *pos += 1;
  }
  step_exit(xa);
}

int launch_argmax(const float* logits, int32_t* tokens_out, int batch, int vocab, cudaStream_t st, const StepState* adv,
                  const ExitArgs& xa) {
  DLI_REQUIRE(logits && tokens_out && batch > 0 && vocab > 0, "argmax: null pointer or empty problem");
  DLI_REQUIRE((reinterpret_cast<uintptr_t>(logits) & 15) == 0 && vocab % 4 == 0, "argmax: logits must be 16-byte aligned rows (vocab %% 4 == 0)");
  const bool a = adv && adv->advance;
  return launch_kernel("argmax_kernel", argmax_kernel, dim3(batch), dim3(ARGMAX_THREADS), 0, st, logits, tokens_out, vocab,
                       a ? adv->seq_lens : (int*)nullptr, a ? adv->pos_rw : (long long*)nullptr, a ? adv->batch : 0, xa);
}

// A stream-ordered wait for a hand-off flag (one thread): what a host-side client puts in front of the copy that reads
// a result another GPU delivers (token ids arriving in stage 0's inbox).
__global__ void flag_wait_kernel(const uint32_t* flag, uint32_t at_least, uint32_t* state, unsigned long long timeout_ns) {
  flag_wait(flag, at_least, state, timeout_ns);
}

int launch_flag_wait(const uint32_t* flag, uint32_t at_least, uint32_t* state, unsigned long long timeout_ns, cudaStream_t st) {
  DLI_REQUIRE(flag && state, "dli_flag_wait: null pointer");
  flag_wait_kernel<<<1, 1, 0, st>>>(flag, at_least, state, timeout_ns);
  DLI_CHECK_LAUNCH("flag_wait_kernel");
  return DLI_OK;
}


// ---------------------------------------------------------------------------------------------------
// Attention-sink eviction (PartialLlamaSinkCache.update, cache.py:111-133, with _get_rerotation_cos_sin
// cache.py:21-48): when n new tokens would reach the window W, the cache becomes
//   [ first `sinks` tokens | re-rotated last (W - sinks - n) keys | (the n new tokens, appended afterwards) ]
// and V is sliced the same way without rotation.  Done in place on the pages, for every layer of the span in one
// launch: CTA = (kv head, sequence, layer) walks its kept tokens left to right in chunks (read chunk, barrier,
// write chunk); destinations always lie at or before sources, so a chunk never overwrites unread data.
// rc = cos[s+n+i] cos[s+i] + sin[s+n+i] sin[s+i],  rs = -sin[s+n+i] cos[s+i] + cos[s+n+i] sin[s+i]   (cache.py:36-41)
// k' = k rc + rotate_half(k) rs                                                (SinkCache._apply_key_rotary_pos_emb)
// cos/sin are the generation's cos/sin history (rows of its first W positions), fp32 [W, hd].
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) kv_evict_kernel(__nv_bfloat16* __restrict__ k_base, __nv_bfloat16* __restrict__ v_base,
                                                       long long layer_stride, const int32_t* __restrict__ block_table,
                                                       int max_pages, int page_tokens, int KV, int hd, int cur_len, int window,
                                                       int sinks, int n_new, const float* __restrict__ cos_hist,
                                                       const float* __restrict__ sin_hist) {
  pdl_trigger();
  pdl_wait();
  const int kvh = blockIdx.x, b = blockIdx.y, layer = blockIdx.z;
  __nv_bfloat16* kp = k_base + (long long)layer * layer_stride;
  __nv_bfloat16* vp = v_base + (long long)layer * layer_stride;
  const int32_t* bt = block_table + (long long)b * max_pages;
  const int keep = window - sinks - n_new;
  const int src0 = cur_len - keep, dst0 = sinks;
  const int half = hd >> 1;
  const int tpt = hd / 16;                 // threads per token: 8 elements of each half per thread
  const int tpi = 256 / tpt;               // tokens per iteration
  const int ti = threadIdx.x / tpt, d0 = (threadIdx.x % tpt) * 8;
  auto addr = [&](int tok) {
    const int page = bt[tok / page_tokens], slot = tok % page_tokens;
    return (((long long)page * KV + kvh) * page_tokens + slot) * hd + d0;
  };
  for (int i0 = 0; i0 < keep; i0 += tpi) {
    const int i = i0 + ti;
    uint4 klo, khi, vlo, vhi;
    uint32_t olo[4], ohi[4];
    if (i < keep) {
      const long long a = addr(src0 + i);
      klo = *reinterpret_cast<const uint4*>(kp + a);
      khi = *reinterpret_cast<const uint4*>(kp + a + half);
      vlo = *reinterpret_cast<const uint4*>(vp + a);
      vhi = *reinterpret_cast<const uint4*>(vp + a + half);
      const float* co = cos_hist + (long long)(sinks + n_new + i) * hd;   // "original" rows
      const float* so = sin_hist + (long long)(sinks + n_new + i) * hd;
      const float* cs = cos_hist + (long long)(sinks + i) * hd;           // "shifted" rows
      const float* ss = sin_hist + (long long)(sinks + i) * hd;
      const uint32_t kl[4] = {klo.x, klo.y, klo.z, klo.w}, kh[4] = {khi.x, khi.y, khi.z, khi.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float lo[2] = {bf16_lo(kl[j]), bf16_hi(kl[j])}, hi[2] = {bf16_lo(kh[j]), bf16_hi(kh[j])}, rl[2], rh[2];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int dl = d0 + 2 * j + e, dh = dl + half;
          const float rc_l = co[dl] * cs[dl] + so[dl] * ss[dl], rs_l = -so[dl] * cs[dl] + co[dl] * ss[dl];
          const float rc_h = co[dh] * cs[dh] + so[dh] * ss[dh], rs_h = -so[dh] * cs[dh] + co[dh] * ss[dh];
          rl[e] = lo[e] * rc_l - hi[e] * rs_l;   // rotate_half: first half pairs with -x2
          rh[e] = hi[e] * rc_h + lo[e] * rs_h;   //              second half pairs with +x1
        }
        olo[j] = pack_bf16(rl[0], rl[1]);
        ohi[j] = pack_bf16(rh[0], rh[1]);
      }
    }
    __syncthreads();  // every source of this chunk is in registers before any destination is written
    if (i < keep) {
      const long long a = addr(dst0 + i);
      *reinterpret_cast<uint4*>(kp + a) = make_uint4(olo[0], olo[1], olo[2], olo[3]);
      *reinterpret_cast<uint4*>(kp + a + half) = make_uint4(ohi[0], ohi[1], ohi[2], ohi[3]);
      *reinterpret_cast<uint4*>(vp + a) = vlo;
      *reinterpret_cast<uint4*>(vp + a + half) = vhi;
    }
    __syncthreads();
  }
}

int launch_kv_evict(const dli_kv_desc* kv, int n_layers, int batch, int KV, int hd, int cur_len, int window, int sinks,
                    int n_new, const float* cos_hist, const float* sin_hist, cudaStream_t st) {
  KvView v;
  int rc = kv_view(kv, 0, &v);
  if (rc) return rc;
  DLI_REQUIRE(cos_hist && sin_hist, "dli_kv_evict: null cos/sin history");
  DLI_REQUIRE(n_layers > 0 && batch > 0 && KV > 0 && (hd == 64 || hd == 128), "dli_kv_evict: bad shape");
  DLI_REQUIRE(sinks >= 0 && n_new > 0 && window - sinks - n_new > 0, "dli_kv_evict: window %d too small for %d sinks + %d new tokens",
              window, sinks, n_new);
  DLI_REQUIRE(cur_len + n_new >= window && cur_len >= window - sinks - n_new,
              "dli_kv_evict: nothing to evict (cur_len=%d n_new=%d window=%d)", cur_len, n_new, window);
  return launch_kernel("kv_evict_kernel", kv_evict_kernel, dim3(KV, batch, n_layers), dim3(256), 0, st, v.k, v.v,
                       (long long)(kv->layer_stride_bytes / 2), v.block_table, v.max_pages, v.page_tokens, KV, hd, cur_len,
                       window, sinks, n_new, cos_hist, sin_hist);
}

DLI_TRACE_SETTER(trace_set_norm)

}  // namespace dli
