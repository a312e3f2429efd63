// Prefill attention over the paged KV cache (q_len >= 1, optional cached past, causal flag).
//
// Replaces modules.py:87-105 for a multi-token chunk: softmax(Q K^T / sqrt(hd) [+ causal mask]) V with GQA by
// indexing.  The [B, H, L, T] score tensor the reference materialises is never formed: flash-attention
// tiling with fp32 online softmax (exp2 domain).  Scores and the running max/sum stay in fp32; P is
// rounded to bf16 only as the tensor-core operand of P.V (the one rounding shared with modules.py:96).
// `causal` == 0 reproduces the reference's default sdpa configuration, which adds no mask at all
// (model.py:94-101 + modules.py:92-94); `causal` == 1 the eager mask of model.py:103-127.
//
// Round-1 implementation: warp-level mma.sync (m16n8k16, bf16 -> fp32) with ldmatrix-fed fragments; one
// CTA = 64 or 128 queries of one head (4 or 8 warps x 16 rows), K/V tiles of 64 keys gathered from pages with
// cp.async into XOR-swizzled shared memory.  Kept for head_dim 64, chunks of fewer than 16 queries and page sizes the
// TMA kernels do not take; head_dim 128 goes to the tcgen05 kernels (attn_prefill_tc.cu, attn_prefill_tc2.cu), which
// this file's launch_prefill_attention dispatches to.
#include "common.cuh"
#include "mma_sync.cuh"

#include <math.h>
#include <stdlib.h>

namespace dli {

constexpr int PF_BN = 64;

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
// byte offset of 16-byte chunk `c` of row `r` in a [rows][HD] bf16 tile, XOR-swizzled against bank conflicts
template <int HD>
__device__ __forceinline__ uint32_t swz(int r, int c) {
  return (uint32_t)(r * HD * 2 + ((c ^ (r & 7)) << 4));
}

// MASKED instantiation (dli_prefill_attention_masked): visibility comes from a per-query key limit and / or a
// per-sequence key validity map instead of the causal flag alone.  Hidden scores take a large FINITE negative value, as
// the reference's additive finfo.min mask does (model.py:106-135): a row with no visible key at all (a left-padding
// query) then comes out as the uniform average of V over the whole context, exactly like the reference's softmax of a
// row of equal entries; rows with at least one visible key give hidden keys probability 0.
constexpr float PF_MASKED = -3.0e38f;

template <int HD, int NWARPS, int NBUF, bool MASKED = false>
__global__ void __launch_bounds__(NWARPS * 32) prefill_attn_kernel(const __nv_bfloat16* __restrict__ q,
                                                                   __nv_bfloat16* __restrict__ out, KvView kv, int q_len,
                                                                   int H, int KV, int causal, float scale_log2,
                                                                   AttnMask mask) {
  constexpr int CPR = HD / 8;          // 16-byte chunks per row
  constexpr int BM = NWARPS * 16;      // queries per CTA
  constexpr int NT = NWARPS * 32;
  constexpr int KV_TILE_BYTES = PF_BN * HD * 2;
  extern __shared__ __align__(128) uint8_t pf_smem[];
  const uint32_t sQ = smem_u32(pf_smem), sK0 = sQ + BM * HD * 2, sV0 = sK0 + NBUF * KV_TILE_BYTES;  // K[NBUF], V[NBUF]

  pdl_trigger();
  pdl_wait();
  const int m0 = blockIdx.x * BM, h = blockIdx.y, b = blockIdx.z;
  const int kvh = h / (H / KV);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t4 = lane & 3;
  const int past = kv.seq_lens[b];
  const int T = past + q_len;
  const int32_t* bt = kv.block_table + (long long)b * kv.max_pages;

  int kv_end = T;
  if (causal && !MASKED) kv_end = min(T, past + m0 + BM);
  const int n_tiles = (kv_end + PF_BN - 1) / PF_BN;

  auto load_kv = [&](int tile) {  // K/V rows of keys [tile*64, +64) -> buffer tile&1 (zeros past T)
    const int j0 = tile * PF_BN;
    const uint32_t sK = sK0 + (tile % NBUF) * KV_TILE_BYTES, sV = sV0 + (tile % NBUF) * KV_TILE_BYTES;
    for (int i = threadIdx.x; i < PF_BN * CPR; i += NT) {
      const int r = i / CPR, c = i - r * CPR;
      const int j = j0 + r;
      const bool ok = j < T;
      long long off = 0;
      if (ok) {
        const int page = bt[j / kv.page_tokens], slot = j % kv.page_tokens;
        off = (((long long)page * KV + kvh) * kv.page_tokens + slot) * HD + c * 8;
      }
      cp_async16(sK + swz<HD>(r, c), kv.k + off, ok ? 16 : 0);
      cp_async16(sV + swz<HD>(r, c), kv.v + off, ok ? 16 : 0);
    }
  };

  // ---- Q tile (rows beyond q_len are zero) + first K/V tile ----
  for (int i = threadIdx.x; i < BM * CPR; i += NT) {
    const int r = i / CPR, c = i - r * CPR;
    const int qi = m0 + r;
    const bool ok = qi < q_len;
    const __nv_bfloat16* src = q + (((long long)b * q_len + (ok ? qi : 0)) * H + h) * HD + c * 8;
    cp_async16(sQ + swz<HD>(r, c), src, ok ? 16 : 0);
  }
  load_kv(0);
  asm volatile("cp.async.commit_group;" ::: "memory");

  uint32_t qa[HD / 16][4];
  float o[HD / 8][4];
#pragma unroll
  for (int n = 0; n < HD / 8; ++n) o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f;
  float mrow[2] = {-INFINITY, -INFINITY}, lrow[2] = {0.f, 0.f};
  const int pos0 = past + m0 + warp * 16 + g;        // absolute position of this thread's first row (second: +8)
  const int warp_last_pos = past + m0 + warp * 16 + 15;

  for (int tile = 0; tile < n_tiles; ++tile) {
    const int j0 = tile * PF_BN;
    if (NBUF == 1 && tile > 0) {  // single buffer (more CTAs per SM hide the load instead)
      load_kv(tile);
      asm volatile("cp.async.commit_group;" ::: "memory");
    }
    if (NBUF == 2 && tile + 1 < n_tiles) {  // prefetch the next tile into the other buffer while this one is consumed
      load_kv(tile + 1);
      asm volatile("cp.async.commit_group;" ::: "memory");
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();
    if (tile == 0) {  // Q fragments of this warp's 16 rows, all k-slices
#pragma unroll
      for (int ks = 0; ks < HD / 16; ++ks) {
        const int r = warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
        const int c = ks * 2 + (lane >> 4);
        ldsm_x4(sQ + swz<HD>(r, c), qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3]);
      }
    }
    const uint32_t sK = sK0 + (tile % NBUF) * KV_TILE_BYTES, sV = sV0 + (tile % NBUF) * KV_TILE_BYTES;

    if (MASKED || !causal || j0 <= warp_last_pos) {  // warp-uniform: tiles entirely above this warp's diagonal are skipped
      // ---- S = Q K^T (16 x 64 per warp) ----
      float s[PF_BN / 8][4];
#pragma unroll
      for (int n = 0; n < PF_BN / 8; ++n) s[n][0] = s[n][1] = s[n][2] = s[n][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < HD / 16; ++ks) {
#pragma unroll
        for (int np = 0; np < PF_BN / 16; ++np) {
          uint32_t b0, b1, b2, b3;
          const int r = np * 16 + (lane & 7) + (lane >> 4) * 8;
          const int c = ks * 2 + ((lane >> 3) & 1);
          ldsm_x4(sK + swz<HD>(r, c), b0, b1, b2, b3);
          mma_bf16_16816(s[2 * np], qa[ks], b0, b1);
          mma_bf16_16816(s[2 * np + 1], qa[ks], b2, b3);
        }
      }

      // ---- mask (only on tiles that touch the diagonal or the end of the context) + online softmax ----
      const bool need_mask = MASKED || (j0 + PF_BN > T) || (causal && j0 + PF_BN - 1 > past + m0 + warp * 16);
      int lim[2] = {T, T};
      if (MASKED && mask.key_limit) {
#pragma unroll
        for (int r2 = 0; r2 < 2; ++r2) {
          const int qi = m0 + warp * 16 + g + r2 * 8;
          lim[r2] = qi < q_len ? mask.key_limit[qi] : 0;
        }
      }
      float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
      for (int n = 0; n < PF_BN / 8; ++n) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          float v = s[n][e] * scale_log2;
          if (need_mask) {
            const int j = j0 + n * 8 + 2 * t4 + (e & 1);
            const int pos = pos0 + (e >> 1) * 8;
            if (MASKED) {
              if (j >= T) v = -INFINITY;
              else if ((mask.key_limit ? j >= lim[e >> 1] : (causal && j > pos)) ||
                       (mask.key_valid && mask.key_valid[(long long)b * mask.key_stride + j] == 0))
                v = PF_MASKED;
            } else if (!(j < T && (!causal || j <= pos))) v = -INFINITY;
          }
          s[n][e] = v;
          mx[e >> 1] = fmaxf(mx[e >> 1], v);
        }
      }
      float corr[2], mref[2];
#pragma unroll
      for (int r2 = 0; r2 < 2; ++r2) {
        mx[r2] = fmaxf(mx[r2], __shfl_xor_sync(0xffffffffu, mx[r2], 1));
        mx[r2] = fmaxf(mx[r2], __shfl_xor_sync(0xffffffffu, mx[r2], 2));
        const float mn = fmaxf(mrow[r2], mx[r2]);
        mref[r2] = mn == -INFINITY ? 0.f : mn;
        corr[r2] = exp2f(mrow[r2] - mref[r2]);
        mrow[r2] = mn;
        lrow[r2] *= corr[r2];
      }
#pragma unroll
      for (int n = 0; n < HD / 8; ++n) {
        o[n][0] *= corr[0]; o[n][1] *= corr[0];
        o[n][2] *= corr[1]; o[n][3] *= corr[1];
      }
      uint32_t pa[PF_BN / 16][4];
#pragma unroll
      for (int n = 0; n < PF_BN / 8; ++n) {
        const float p0 = exp2f(s[n][0] - mref[0]), p1 = exp2f(s[n][1] - mref[0]);
        const float p2 = exp2f(s[n][2] - mref[1]), p3 = exp2f(s[n][3] - mref[1]);
        lrow[0] += p0 + p1;
        lrow[1] += p2 + p3;
        // C layout of two adjacent 8-key tiles == A layout of one 16-key slice
        pa[n >> 1][(n & 1) * 2 + 0] = pack_bf16(p0, p1);
        pa[n >> 1][(n & 1) * 2 + 1] = pack_bf16(p2, p3);
      }

      // ---- O += P V ----
#pragma unroll
      for (int ks = 0; ks < PF_BN / 16; ++ks) {
#pragma unroll
        for (int np = 0; np < HD / 16; ++np) {
          uint32_t b0, b1, b2, b3;
          const int r = ks * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
          const int c = np * 2 + (lane >> 4);
          ldsm_x4_trans(sV + swz<HD>(r, c), b0, b1, b2, b3);
          mma_bf16_16816(o[2 * np], pa[ks], b0, b1);
          mma_bf16_16816(o[2 * np + 1], pa[ks], b2, b3);
        }
      }
    }
    __syncthreads();  // every warp is done with this buffer before the load of tile + 2 overwrites it
  }

  // ---- finalize: row sums across the quad, normalise, store ----
#pragma unroll
  for (int r2 = 0; r2 < 2; ++r2) {
    lrow[r2] += __shfl_xor_sync(0xffffffffu, lrow[r2], 1);
    lrow[r2] += __shfl_xor_sync(0xffffffffu, lrow[r2], 2);
  }
#pragma unroll
  for (int r2 = 0; r2 < 2; ++r2) {
    const int qi = m0 + warp * 16 + g + r2 * 8;
    if (qi < q_len) {
      const float inv = lrow[r2] > 0.f ? 1.f / lrow[r2] : 0.f;
      __nv_bfloat16* dst = out + (((long long)b * q_len + qi) * H + h) * HD + 2 * t4;
#pragma unroll
      for (int n = 0; n < HD / 8; ++n)
        *reinterpret_cast<uint32_t*>(dst + n * 8) = pack_bf16(o[n][r2 * 2] * inv, o[n][r2 * 2 + 1] * inv);
    }
  }
}

template <int HD, int NWARPS, int NBUF>
static int launch_pf(const __nv_bfloat16* q, __nv_bfloat16* out, const KvView& kv, int batch, int q_len, int H, int KV,
                     bool causal, cudaStream_t st) {
  constexpr int BM = NWARPS * 16;
  const size_t smem = (size_t)(BM + 2 * NBUF * PF_BN) * HD * 2;
  DLI_ONCE_PER_DEVICE(cudaFuncSetAttribute(prefill_attn_kernel<HD, NWARPS, NBUF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));  // > 48 KB dynamic shared memory: opt-in per device
  const float scale_log2 = 1.4426950408889634f / sqrtf((float)HD);
  dim3 grid((q_len + BM - 1) / BM, H, batch);
  return launch_kernel("prefill_attn_kernel", prefill_attn_kernel<HD, NWARPS, NBUF>, grid, dim3(NWARPS * 32), smem, st, q, out, kv,
                       q_len, H, KV, causal ? 1 : 0, scale_log2, AttnMask{nullptr, nullptr, 0});
}

template <int HD>
static int launch_pf_masked(const __nv_bfloat16* q, __nv_bfloat16* out, const KvView& kv, int batch, int q_len, int H, int KV,
                            bool causal, const AttnMask& mask, cudaStream_t st) {
  constexpr int NWARPS = 4, NBUF = 1, BM = NWARPS * 16;
  const size_t smem = (size_t)(BM + 2 * NBUF * PF_BN) * HD * 2;
  DLI_ONCE_PER_DEVICE(cudaFuncSetAttribute(prefill_attn_kernel<HD, NWARPS, NBUF, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const float scale_log2 = 1.4426950408889634f / sqrtf((float)HD);
  dim3 grid((q_len + BM - 1) / BM, H, batch);
  return launch_kernel("prefill_attn_kernel<masked>", prefill_attn_kernel<HD, NWARPS, NBUF, true>, grid, dim3(NWARPS * 32), smem, st,
                       q, out, kv, q_len, H, KV, causal ? 1 : 0, scale_log2, mask);
}

int launch_prefill_attention_masked(const __nv_bfloat16* q, __nv_bfloat16* out, const KvView& kv, int batch, int q_len, int H,
                                    int KV, int hd, bool causal, const AttnMask& mask, cudaStream_t st) {
  DLI_REQUIRE(q && out, "dli_prefill_attention_masked: null pointer");
  DLI_REQUIRE(batch > 0 && q_len > 0 && H > 0 && KV > 0 && H % KV == 0, "dli_prefill_attention_masked: batch=%d q_len=%d H=%d KV=%d",
              batch, q_len, H, KV);
  DLI_REQUIRE(!mask.key_valid || mask.key_stride >= kv.max_ctx, "dli_prefill_attention_masked: key_valid rows of %d entries < context %d",
              mask.key_stride, kv.max_ctx);
  if (hd == 128) return launch_pf_masked<128>(q, out, kv, batch, q_len, H, KV, causal, mask, st);
  if (hd == 64) return launch_pf_masked<64>(q, out, kv, batch, q_len, H, KV, causal, mask, st);
  return fail(DLI_ERR_UNSUPPORTED, "dli_prefill_attention_masked: head_dim %d is not instantiated (64, 128)", hd);
}

int launch_prefill_attention(const __nv_bfloat16* q, __nv_bfloat16* out, const KvView& kv, int batch, int q_len, int H,
                             int KV, int hd, bool causal, cudaStream_t st) {
  DLI_REQUIRE(q && out, "dli_prefill_attention: null pointer");
  DLI_REQUIRE(batch > 0 && q_len > 0 && H > 0 && KV > 0 && H % KV == 0, "dli_prefill_attention: batch=%d q_len=%d H=%d KV=%d",
              batch, q_len, H, KV);
    // head_dim 128: both contractions on tcgen05 (attn_prefill_tc.cu); q_len >= 16 keeps tiny chunks on the light kernel
  if (prefill_attn_tc2_supported(kv, hd, q_len)) return launch_prefill_attention_tc2(q, out, kv, batch, q_len, H, KV, causal, st);
  if (q_len >= 16 && prefill_attn_tc_supported(kv, hd)) return launch_prefill_attention_tc(q, out, kv, batch, q_len, H, KV, causal, st);
  static int force = -1;
  if (force < 0) {
    const char* e = exp_env("DLI_PF_WARPS");
    force = e ? atoi(e) : 0;
  }
  // measured on B200 (70B shape, 8k tokens): 4 warps + single K/V buffer (48 KB, 3 CTAs/SM) 68.7 ms per 10 layers,
  // 4 warps double-buffered 77.2 ms, 8 warps double-buffered 75.7 ms: occupancy beats intra-CTA pipelining here
  if (force == 8) {
    if (hd == 128) return launch_pf<128, 8, 2>(q, out, kv, batch, q_len, H, KV, causal, st);
    if (hd == 64) return launch_pf<64, 8, 2>(q, out, kv, batch, q_len, H, KV, causal, st);
  }
  if (hd == 128) return launch_pf<128, 4, 1>(q, out, kv, batch, q_len, H, KV, causal, st);
  if (hd == 64) return launch_pf<64, 4, 1>(q, out, kv, batch, q_len, H, KV, causal, st);
  return fail(DLI_ERR_UNSUPPORTED, "dli_prefill_attention: head_dim %d is not instantiated (64, 128)", hd);
}

}  // namespace dli
