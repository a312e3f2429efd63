// RoPE applied where the rotated value is consumed (decode attention with an unsplit context): the q|k|v projection
// of the step's single token is still in the fp32 split-K accumulation buffer, and the one CTA that owns a
// (sequence, KV head) rotates its query heads and the new key (modules.py:17-20: x*cos + rotate_half(x)*sin with
// rotate_half = cat(-x2, x1)), appends key and value to the page (cache.py:103-109) and re-zeroes what it consumed —
// the work of rope_kv_append_kernel without its launch.
#pragma once
#include "common.cuh"

namespace dli {

// rotated dims [d0, d0 + 8) of one head whose fp32 projection starts at x; d0 is a multiple of 8
template <int HD>
__device__ __forceinline__ void rope8(const float* __restrict__ x, const float* __restrict__ cs,
                                      const float* __restrict__ sn, int d0, float (&out)[8]) {
  constexpr int HALF = HD / 2;
  const int dp = d0 < HALF ? d0 + HALF : d0 - HALF;
  const float sign = d0 < HALF ? -1.f : 1.f;
  const float4 a0 = *reinterpret_cast<const float4*>(x + d0), a1 = *reinterpret_cast<const float4*>(x + d0 + 4);
  const float4 p0 = *reinterpret_cast<const float4*>(x + dp), p1 = *reinterpret_cast<const float4*>(x + dp + 4);
  const float4 c0 = *reinterpret_cast<const float4*>(cs + d0), c1 = *reinterpret_cast<const float4*>(cs + d0 + 4);
  const float4 s0 = *reinterpret_cast<const float4*>(sn + d0), s1 = *reinterpret_cast<const float4*>(sn + d0 + 4);
  out[0] = a0.x * c0.x + sign * p0.x * s0.x; out[1] = a0.y * c0.y + sign * p0.y * s0.y;
  out[2] = a0.z * c0.z + sign * p0.z * s0.z; out[3] = a0.w * c0.w + sign * p0.w * s0.w;
  out[4] = a1.x * c1.x + sign * p1.x * s1.x; out[5] = a1.y * c1.y + sign * p1.y * s1.y;
  out[6] = a1.z * c1.z + sign * p1.z * s1.z; out[7] = a1.w * c1.w + sign * p1.w * s1.w;
}

// rotated dims (d, d + 1), d even, packed as two bf16
template <int HD>
__device__ __forceinline__ uint32_t rope2_bf16(const float* __restrict__ x, const float* __restrict__ cs,
                                               const float* __restrict__ sn, int d) {
  constexpr int HALF = HD / 2;
  const int dp = d < HALF ? d + HALF : d - HALF;
  const float sign = d < HALF ? -1.f : 1.f;
  const float2 a = *reinterpret_cast<const float2*>(x + d), p = *reinterpret_cast<const float2*>(x + dp);
  const float2 c = *reinterpret_cast<const float2*>(cs + d), s = *reinterpret_cast<const float2*>(sn + d);
  return pack_bf16(a.x * c.x + sign * p.x * s.x, a.y * c.y + sign * p.y * s.y);
}

// The step's own key (rotated) and value rows as the pages store them: lanes [0, HD/8) return the 16-byte chunk
// `lane` of the key row, lanes [HD/8, HD/4) that of the value row; other lanes return zeros.
// cos/sin rows of the step: the pointers themselves, or row *pos of the tables they point at
template <int HD>
__device__ __forceinline__ DecRope at_step_row(DecRope rp) {
  if (rp.pos) {
    const long long r = *rp.pos;
    rp.cos += r * HD;
    rp.sin += r * HD;
  }
  return rp;
}

template <int HD>
__device__ __forceinline__ uint4 new_kv_chunk(const DecRope& rp, int b, int H, int KV, int kvh, int lane) {
  constexpr int CPR = HD / 8;
  uint4 r = make_uint4(0, 0, 0, 0);
  const float* row = rp.qkv + (long long)b * rp.n_all;
  if (lane < CPR) {
    float k[8];
    rope8<HD>(row + (H + kvh) * HD, rp.cos, rp.sin, lane * 8, k);
    r = make_uint4(pack_bf16(k[0], k[1]), pack_bf16(k[2], k[3]), pack_bf16(k[4], k[5]), pack_bf16(k[6], k[7]));
  } else if (lane < 2 * CPR) {
    const float* v = row + (H + KV + kvh) * HD + (lane - CPR) * 8;
    const float4 a = *reinterpret_cast<const float4*>(v), c = *reinterpret_cast<const float4*>(v + 4);
    r = make_uint4(pack_bf16(a.x, a.y), pack_bf16(a.z, a.w), pack_bf16(c.x, c.y), pack_bf16(c.z, c.w));
  }
  return r;
}

// store that chunk to the page slot of position `pos` (one warp calls this once per (sequence, KV head))
template <int HD>
__device__ __forceinline__ void append_new_kv(const KvView& kv, int b, int KV, int kvh, int pos, int lane, uint4 chunk) {
  constexpr int CPR = HD / 8;
  if (lane >= 2 * CPR) return;
  const int page = kv.block_table[(long long)b * kv.max_pages + pos / kv.page_tokens], slot = pos % kv.page_tokens;
  __nv_bfloat16* base = (lane < CPR ? kv.k : kv.v) + (((long long)page * KV + kvh) * kv.page_tokens + slot) * HD;
  *reinterpret_cast<uint4*>(base + (lane % CPR) * 8) = chunk;
}

// re-zero the (G + 2) heads of the accumulation row this CTA consumed
template <int HD, int G, int NTHREADS>
__device__ __forceinline__ void zero_consumed(const DecRope& rp, int b, int H, int KV, int kvh) {
  float* row = const_cast<float*>(rp.qkv) + (long long)b * rp.n_all;
  for (int i = threadIdx.x * 4; i < (G + 2) * HD; i += NTHREADS * 4) {
    const int head = i / HD, d = i - head * HD;
    float* dst = head < G ? row + (kvh * G + head) * HD + d
                          : head == G ? row + (H + kvh) * HD + d : row + (H + KV + kvh) * HD + d;
    *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

}  // namespace dli
