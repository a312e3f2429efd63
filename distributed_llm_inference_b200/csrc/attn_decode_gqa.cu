// Flash-decode attention for grouped-query models (q_len == 1, H / KV in {2, 4, 8}).
//
// Replaces modules.py:87-105 for a decode step of a GQA model (Llama-2-70B: 64 query heads on 8 KV heads).
// The step is HBM-bound only if the arithmetic keeps up with the K/V stream: with 8 query heads per KV head
// the CUDA-core kernel in attn_decode.cu spends 16 FMAs and a shuffle tree per loaded bf16 pair and ran at
// 0.9-1.1 TB/s (ncu, 70B shapes).  Here the G query heads of a group are the rows of an m16n8k16 tile
// (rows >= G are zero padding), so one warp-level MMA covers what took 128 FMAs:
//
//   S[16 x 16 tokens] = Q[16 x hd] . K^T        (K rows straight from the paged pool, ldmatrix)
//   O[16 x hd]       += P[16 x 16 tokens] . V    (P from the S registers, V through ldmatrix.trans)
//
// grid = (splits, KV heads, batch) as in attn_decode.cu, same split-K partials and the same combine kernel.
// A CTA is 4 consumer warps + 1 producer warp.  The producer walks the block table and pulls whole pages
// with 2-D TMA ([page_tokens x 64] boxes, 128-byte swizzle, so ldmatrix is bank-conflict free) into a ring
// of GQ_STAGES x 64-token stages; every consumer warp owns 16 tokens of a stage and keeps its own online
// softmax state (fp32, exp2 domain), merged through shared memory at the end.  P is rounded to bf16 only as
// the tensor-core operand of P.V -- the rounding modules.py:96 applies before its own matmul.
#include "common.cuh"
#include "mma_sync.cuh"
#include "tc_ptx.cuh"
#include "rope_inline.cuh"

#include <math.h>
#include <stdlib.h>

namespace dli {

constexpr int GQ_STAGE_TOK = 64;
constexpr int GQ_STAGES = 3;
constexpr int GQ_WARPS = 4;
constexpr int GQ_THREADS = (GQ_WARPS + 1) * 32;

template <int HD>
__host__ __device__ constexpr int gq_stage_bytes() { return GQ_STAGE_TOK * HD * 2 * 2; }  // K then V
template <int HD, int G>
__host__ __device__ constexpr int gq_smem_bytes() {
  constexpr int ring = GQ_STAGES * gq_stage_bytes<HD>() + 2 * GQ_STAGES * 8;
  constexpr int merge = GQ_WARPS * G * (HD + 2) * 4;
  return (ring > merge ? ring : merge) + 1024;  // + slack to align the ring to the 1 KB swizzle atom
}

// byte offset of 16-byte chunk `ch` of token row `row` inside a K (or V) stage part: [hd / 64][64 tokens][128 B],
// chunks XOR-ed with the row as the TMA 128-byte swizzle stores them
__device__ __forceinline__ uint32_t gq_off(int row, int ch) {
  return (uint32_t)((ch >> 3) * (GQ_STAGE_TOK * 128) + row * 128 + (((ch & 7) ^ (row & 7)) << 4));
}

template <int HD, int G>
__global__ void __launch_bounds__(GQ_THREADS) decode_attn_gqa_kernel(const __grid_constant__ CUtensorMap tk,
                                                                     const __grid_constant__ CUtensorMap tv,
                                                                     const __nv_bfloat16* __restrict__ q, KvView kv,
                                                                     float* __restrict__ scratch,
                                                                     __nv_bfloat16* __restrict__ out, int H, int KV,
                                                                     float scale_log2, int early_kv,
                                                                     const DecRope rp_in) {
  static_assert(G <= 8, "one m16 tile holds at most 8 real heads in its lower half");
  constexpr int HALVES = HD / 64;
  constexpr int HALF_BYTES = GQ_STAGE_TOK * 128;
  constexpr int KPART = HALVES * HALF_BYTES;
  constexpr int STAGE = gq_stage_bytes<HD>();
  constexpr int KS = HD / 16;  // k-steps of Q.K^T
  constexpr int NT = HD / 8;   // n-tiles of P.V
  constexpr int CPR = HD / 8;  // 16-byte chunks per token row

  extern __shared__ unsigned char dsm_raw[];
  const uint32_t ring = (smem_u32(dsm_raw) + 1023u) & ~1023u;
  unsigned char* dsm = dsm_raw + (ring - smem_u32(dsm_raw));
  const uint32_t full0 = ring + GQ_STAGES * STAGE, empty0 = full0 + 8 * GQ_STAGES;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  if (threadIdx.x == 0) {
    for (int i = 0; i < GQ_STAGES; ++i) {
      mbar_init(full0 + 8 * i, 1);
      mbar_init(empty0 + 8 * i, GQ_WARPS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  pdl_trigger();
  if (threadIdx.x == 0) DLI_TRACE_EV(TR_ATTN_GQA, 0);
  // early_kv (see common.cuh): the producer warp streams the pages of past tokens before the dependency wait
  const bool producer = warp == GQ_WARPS;
  if (!(producer && early_kv)) pdl_wait();
  if (threadIdx.x == 0) DLI_TRACE_EV(TR_ATTN_GQA, 1);

  const int split = blockIdx.x, n_splits = gridDim.x, kvh = blockIdx.y, b = blockIdx.z;
  const int T = kv.seq_lens[b] + 1;  // the step's own token is (being) appended at position T - 1
  int chunk = (T + n_splits - 1) / n_splits;
  chunk = (chunk + GQ_STAGE_TOK - 1) & ~(GQ_STAGE_TOK - 1);
  const int t_begin = split * chunk;
  const int t_end = min(T, t_begin + chunk);
  const int n_stages = t_end > t_begin ? (t_end - t_begin + GQ_STAGE_TOK - 1) / GQ_STAGE_TOK : 0;
  // fused RoPE + append (rope_inline.cuh): the pages hold tokens [0, T - 1); the step's own key and value go to the
  // page and straight into their row of the last ring stage.  Split context: every split rotates its own copy of the
  // queries, the one split whose range holds position T - 1 appends, the combine kernel re-zeroes the projection buffer.
  const DecRope rp = at_step_row<HD>(rp_in);
  const bool fused = rp.qkv != nullptr;
  const bool owns_new = fused && T - 1 >= t_begin && T - 1 < t_end;
  const int t_copy = owns_new ? T - 1 : T;

  const int r = lane >> 2, c = lane & 3;  // fragment row (query head of the group) and column pair
  float o[NT][4];
#pragma unroll
  for (int j = 0; j < NT; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
  float m_run = -INFINITY, l_run = 0.f;  // row r; l is this lane's share, reduced over the quad at the end

  if (warp == GQ_WARPS) {
    // ===== producer warp: lane i pulls the i-th page of the stage (K and V, hd / 64 boxes each) =====
    const int32_t* bt = kv.block_table + (long long)b * kv.max_pages;
    const int pt = kv.page_tokens;
    bool waited = !early_kv;
    for (int i = 0; i < n_stages; ++i) {
      const int slot = i % GQ_STAGES;
      const int s0 = t_begin + i * GQ_STAGE_TOK, s1 = min(min(t_end, t_copy), s0 + GQ_STAGE_TOK);
      const int n_pages = s1 > s0 ? (s1 - s0 + pt - 1) / pt : 0;  // only pages that hold a (past) token of this split
      int page = -1;
      if (lane < n_pages) page = bt[s0 / pt + lane];
      if (!waited && s1 >= T) {  // this stage holds the step's own token: not before the append kernel is done
        pdl_wait();
        waited = true;
      }
      if (i >= GQ_STAGES) mbar_wait(empty0 + 8 * slot, ((uint32_t)(i / GQ_STAGES) - 1) & 1);
      if (lane == 0) mbar_expect_tx(full0 + 8 * slot, (uint32_t)n_pages * pt * HD * 2 * 2);
      __syncwarp();
      if (page >= 0) {
        const int row = (page * KV + kvh) * pt;
        const uint32_t dst = ring + slot * STAGE + (uint32_t)(lane * pt) * 128;
#pragma unroll
        for (int h = 0; h < HALVES; ++h) {
          tma_load_2d(dst + h * HALF_BYTES, &tk, full0 + 8 * slot, h * 64, row);
          tma_load_2d(dst + KPART + h * HALF_BYTES, &tv, full0 + 8 * slot, h * 64, row);
        }
      }
    }
    if (!waited) pdl_wait();  // this warp also stores results below
  } else {
    // Q fragments of the group: rows r < G are real heads, the rest of the m16 tile is zero
    uint32_t qa[KS][2];
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      qa[ks][0] = qa[ks][1] = 0u;
      if (r < G) {
        if (fused) {
          const float* x = rp.qkv + (long long)b * rp.n_all + (kvh * G + r) * HD;
          qa[ks][0] = rope2_bf16<HD>(x, rp.cos, rp.sin, ks * 16 + 2 * c);
          qa[ks][1] = rope2_bf16<HD>(x, rp.cos, rp.sin, ks * 16 + 2 * c + 8);
        } else {
          const __nv_bfloat16* src = q + ((long long)b * H + kvh * G + r) * HD + ks * 16 + 2 * c;
          qa[ks][0] = *reinterpret_cast<const uint32_t*>(src);
          qa[ks][1] = *reinterpret_cast<const uint32_t*>(src + 8);
        }
      }
    }
    const int tw = warp * 16;  // this warp's token rows inside a stage
    const int new_row = (T - 1 - t_begin) % GQ_STAGE_TOK;  // row of the step's token inside the last stage
    uint4 newrow = make_uint4(0, 0, 0, 0);
    if (owns_new) {
      newrow = new_kv_chunk<HD>(rp, b, H, KV, kvh, lane);
      if (warp == 0) append_new_kv<HD>(kv, b, KV, kvh, T - 1, lane, newrow);
    }
    for (int i = 0; i < n_stages; ++i) {
      const int slot = i % GQ_STAGES;
      const int s0 = t_begin + i * GQ_STAGE_TOK;
      const uint32_t kbase = ring + slot * STAGE, vbase = kbase + KPART;
      mbar_wait(full0 + 8 * slot, (uint32_t)(i / GQ_STAGES) & 1);
      if (owns_new && i == n_stages - 1 && warp == new_row / 16) {
        // the warp that owns that token row puts the new K and V there (TMA wrote at most a stale copy of it)
        if (lane < 2 * CPR) {
          const uint32_t dst = (lane < CPR ? kbase : vbase) + gq_off(new_row, lane % CPR);
          asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(newrow.x), "r"(newrow.y), "r"(newrow.z),
                       "r"(newrow.w)
                       : "memory");
        }
        __syncwarp();
      }
      const int valid = t_end - s0;  // token rows >= valid hold whatever was there before
      if (valid < tw + 16) {
        // P is zero for those rows, but 0 * NaN is NaN: clear the V rows this warp is about to multiply
        for (int idx = lane; idx < 16 * CPR; idx += 32) {
          const int row = tw + idx / CPR, ch = idx % CPR;
          if (row >= valid)
            asm volatile("st.shared.v4.u32 [%0], {%1, %1, %1, %1};" ::"r"(vbase + gq_off(row, ch)), "r"(0u) : "memory");
        }
        __syncwarp();
      }

      // ---- S = Q . K^T for 16 tokens (two n-tiles of 8)
      float s[2][4];
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
#pragma unroll
      for (int k2 = 0; k2 < KS / 2; ++k2) {
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
          uint32_t b0, b1, b2, b3;
          ldsm_x4(kbase + gq_off(tw + nt * 8 + (lane & 7), 4 * k2 + (lane >> 3)), b0, b1, b2, b3);
          const uint32_t a0[4] = {qa[2 * k2][0], 0u, qa[2 * k2][1], 0u};
          const uint32_t a1[4] = {qa[2 * k2 + 1][0], 0u, qa[2 * k2 + 1][1], 0u};
          mma_bf16_16816(s[nt], a0, b0, b1);
          mma_bf16_16816(s[nt], a1, b2, b3);
        }
      }

      // ---- online softmax of row r over the 4 scores this lane holds (tokens tw + nt*8 + 2c + e)
      float mx = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const bool ok = tw + nt * 8 + 2 * c + e < valid;
          s[nt][e] = ok ? s[nt][e] * scale_log2 : -INFINITY;
          mx = fmaxf(mx, s[nt][e]);
        }
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      const float mn = fmaxf(m_run, mx);
      const float mref = mn == -INFINITY ? 0.f : mn;  // nothing seen yet: keep everything at zero
      const float corr = exp2f(m_run - mref);
      m_run = mn;
      float p[2][2];
      float psum = 0.f;
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          p[nt][e] = exp2f(s[nt][e] - mref);
          psum += p[nt][e];
        }
      l_run = l_run * corr + psum;
      const uint32_t pa[4] = {pack_bf16(p[0][0], p[0][1]), 0u, pack_bf16(p[1][0], p[1][1]), 0u};

      // ---- O = O * corr + P . V
#pragma unroll
      for (int j = 0; j < NT; ++j) { o[j][0] *= corr; o[j][1] *= corr; }
#pragma unroll
      for (int j2 = 0; j2 < NT / 2; ++j2) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4_trans(vbase + gq_off(tw + ((lane >> 3) & 1) * 8 + (lane & 7), 2 * j2 + (lane >> 4)), b0, b1, b2, b3);
        mma_bf16_16816(o[2 * j2], pa, b0, b1);
        mma_bf16_16816(o[2 * j2 + 1], pa, b2, b3);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty0 + 8 * slot);
    }
    l_run += __shfl_xor_sync(0xffffffffu, l_run, 1);
    l_run += __shfl_xor_sync(0xffffffffu, l_run, 2);
  }

  // merge the 4 warps through the (now idle) ring
  __syncthreads();
  float* sm_o = reinterpret_cast<float*>(dsm);  // [GQ_WARPS][G][HD]
  float* sm_m = sm_o + GQ_WARPS * G * HD;       // [GQ_WARPS][G]
  float* sm_l = sm_m + GQ_WARPS * G;
  if (warp < GQ_WARPS && r < G) {
#pragma unroll
    for (int j = 0; j < NT; ++j)
      *reinterpret_cast<float2*>(&sm_o[(warp * G + r) * HD + 8 * j + 2 * c]) = make_float2(o[j][0], o[j][1]);
    if (c == 0) { sm_m[warp * G + r] = m_run; sm_l[warp * G + r] = l_run; }
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < G * HD; idx += GQ_THREADS) {
    const int g = idx / HD, d = idx - g * HD;
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < GQ_WARPS; ++w) M = fmaxf(M, sm_m[w * G + g]);
    float L = 0.f, O = 0.f;
    if (M > -INFINITY) {
#pragma unroll
      for (int w = 0; w < GQ_WARPS; ++w) {
        const float wt = exp2f(sm_m[w * G + g] - M);  // warps that saw no token have m = -inf -> weight 0
        L = fmaf(sm_l[w * G + g], wt, L);
        O = fmaf(sm_o[(w * G + g) * HD + d], wt, O);
      }
    }
    if (n_splits == 1) {
      out[((long long)b * H + kvh * G + g) * HD + d] = __float2bfloat16_rn(O / L);
    } else {
      float* dst = scratch + (((long long)b * H + kvh * G + g) * n_splits + split) * (HD + 2);
      dst[d] = O;
      if (d == 0) { dst[HD] = M; dst[HD + 1] = L; }
    }
  }
  if (fused && n_splits == 1) zero_consumed<HD, G, GQ_THREADS>(rp, b, H, KV, kvh);  // every read of the row is behind the barriers above
  if (threadIdx.x == 0) DLI_TRACE_EV(TR_ATTN_GQA, 2);
}

bool decode_attn_gqa_supported(const KvView& kv, int G, int hd) {
  static const bool off = exp_env("DLI_DEC_DIRECT") != nullptr || exp_env("DLI_DEC_NO_MMA") != nullptr;
  if (off) return false;
  if (!(G == 2 || G == 4 || G == 8) || !(hd == 64 || hd == 128)) return false;
  const int pt = kv.page_tokens;
  return pt >= 8 && pt % 8 == 0 && GQ_STAGE_TOK % pt == 0;  // whole pages per stage, 1 KB swizzle atoms intact
}

template <int HD, int G>
static int launch_gqa(const CUtensorMap& tk, const CUtensorMap& tv, const __nv_bfloat16* q, __nv_bfloat16* out,
                      const KvView& kv, int batch, int H, int KV, int splits, float* scratch, bool early_kv,
                      const DecRope& rope, cudaStream_t st) {
  DLI_ONCE_PER_DEVICE(cudaFuncSetAttribute(decode_attn_gqa_kernel<HD, G>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     gq_smem_bytes<HD, G>()));  // > 48 KB dynamic shared memory: opt-in per device
  const float scale_log2 = 1.4426950408889634f / sqrtf((float)HD);
  return launch_kernel("decode_attn_gqa_kernel", decode_attn_gqa_kernel<HD, G>, dim3(splits, KV, batch), dim3(GQ_THREADS),
                       (size_t)gq_smem_bytes<HD, G>(), st, tk, tv, q, kv, scratch, out, H, KV, scale_log2, early_kv ? 1 : 0, rope);
}

// Launches the main kernel only; the caller runs the shared combine kernel when splits > 1.
int launch_decode_attention_gqa(const __nv_bfloat16* q, __nv_bfloat16* out, const KvView& kv, int batch, int H, int KV,
                                int hd, int splits, float* scratch, bool early_kv, const DecRope& rope, cudaStream_t st) {
  const int G = H / KV;
  CUtensorMap tk, tv;
  const long long pool_rows = (long long)kv.n_pages * KV * kv.page_tokens;
  int rc = make_tmap_bf16(&tk, kv.k, pool_rows, hd, hd, kv.page_tokens);
  if (rc) return rc;
  if ((rc = make_tmap_bf16(&tv, kv.v, pool_rows, hd, hd, kv.page_tokens))) return rc;
#define DLI_GQA_CASE(HD_, G_) \
  if (hd == HD_ && G == G_) return launch_gqa<HD_, G_>(tk, tv, q, out, kv, batch, H, KV, splits, scratch, early_kv, rope, st)
  DLI_GQA_CASE(128, 2); DLI_GQA_CASE(128, 4); DLI_GQA_CASE(128, 8);
  DLI_GQA_CASE(64, 2); DLI_GQA_CASE(64, 4); DLI_GQA_CASE(64, 8);
#undef DLI_GQA_CASE
  return fail(DLI_ERR_UNSUPPORTED, "dli_decode_attention: GQA kernel not instantiated for head_dim %d group %d", hd, G);
}

DLI_TRACE_SETTER(trace_set_attn_gqa)

}  // namespace dli
