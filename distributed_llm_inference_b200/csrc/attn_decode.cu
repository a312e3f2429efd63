// Flash-decode attention over the paged KV cache (q_len == 1), multi-head models and the generic fallbacks.
//
// Replaces modules.py:87-105 for a decode step: repeat_kv + QK^T/sqrt(hd) + fp32 softmax + PV + head
// merge.  HBM-bound (every cached K and V element is read exactly once per step).
//
// grid = (splits, KV heads, batch).  A CTA owns one (sequence, kv head, context split).  Three kernels:
//   decode_attn_kernel         (default) a producer warp streams pages into a shared-memory ring with 1-D bulk
//                              copies; 4 consumer warps, cut into token groups of hd/8 lanes that each hold an 8-wide
//                              slice of head_dim, do the dot products (warp-shuffle reduce) and the online softmax;
//   decode_attn_direct_kernel  the first version (128-bit global loads, register double buffering), DLI_DEC_DIRECT=1;
//   decode_attn_gqa_kernel     (attn_decode_gqa.cu) grouped-query models on tensor cores.
// Scores, running max, running sum and the output accumulator stay in fp32 (exp2 domain).  A split context leaves
// per-split partials (unnormalised o, m, l) in caller-provided fp32 scratch and decode_combine_kernel merges them; an
// unsplit one is normalised and stored directly and, inside the span, also does the RoPE/append kernel's work
// (rope_inline.cuh).  This file also holds the split heuristic and the allocation-free entry point.
#include "common.cuh"

#include <math.h>
#include <stdlib.h>

#include "tc_ptx.cuh"
#include "rope_inline.cuh"

namespace dli {

constexpr int DEC_THREADS = 128;
constexpr int DEC_MAX_SPLITS = 32;

// Context splits per (sequence, KV head): as many as keep the whole grid RESIDENT AT ONCE (`slots` = co-resident CTAs
// of the kernel that will run), never more.  The kernels are HBM-bound and every CTA keeps its own ring of K/V in
// flight, so one full wave streams at full rate however the work is cut; a second, partly filled wave does not (ncu,
// 70B B=32: 512 CTAs on 296 slots ran 1.73 waves at 3.2 TB/s), and splits == 1 also spares the combine launch.
static int decode_splits(int batch, int KV, int max_ctx, int slots) {
  const int by_ctx = (max_ctx + 63) / 64;
  const int pairs = batch * KV;
  int s = pairs >= slots ? 1 : slots / pairs;
  if (s > by_ctx) s = by_ctx;
  if (s > DEC_MAX_SPLITS) s = DEC_MAX_SPLITS;
  return s < 1 ? 1 : s;
}

// the split count launch_decode_attention will use for this problem (also sizes the static-scratch check)
static int decode_splits_for(const KvView& kv, int batch, int H, int KV, int hd) {
  const int G = H / KV;
  const bool gqa = decode_attn_gqa_supported(kv, G, hd);
  const bool direct_env = exp_env("DLI_DEC_DIRECT") != nullptr;
  // co-resident CTAs per SM: 2 for the tensor-core GQA kernel (96 KB ring) and the register-staged fallback, 4 for
  // the staged kernel (48 KB ring, 70 registers) as long as its register-file share allows (G <= 2)
  const int per_sm = gqa ? 2 : (direct_env || G > 2) ? 2 : 4;
  return decode_splits(batch, KV, kv.max_ctx, per_sm * sm_count());
}

size_t decode_attn_scratch_bytes(int batch, int H, int hd, int max_ctx) {
  int s = (max_ctx + 63) / 64;
  if (s > DEC_MAX_SPLITS) s = DEC_MAX_SPLITS;
  if (s < 1) s = 1;
  return (size_t)batch * H * s * (hd + 2) * sizeof(float);
}

template <int HD, int G>
__global__ void __launch_bounds__(DEC_THREADS) decode_attn_direct_kernel(const __nv_bfloat16* __restrict__ q, KvView kv,
                                                                  float* __restrict__ scratch,
                                                                  __nv_bfloat16* __restrict__ out, int H, int KV,
                                                                  float scale_log2, AttnMask mask) {
  constexpr int LPT = HD / 8;                // lanes per token
  constexpr int GPW = 32 / LPT;              // token groups per warp
  constexpr int NG = (DEC_THREADS / 32) * GPW;  // token groups per CTA
  constexpr int UNROLL = G <= 2 ? 4 : 2;

  pdl_trigger();
  pdl_wait();
  const int split = blockIdx.x, n_splits = gridDim.x, kvh = blockIdx.y, b = blockIdx.z;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int grp = warp * GPW + lane / LPT;   // token group of this lane
  const int dl = (lane % LPT) * 8;           // head_dim slice [dl, dl + 8)

  const int T = kv.seq_lens[b] + 1;          // the step's own token is already appended
  int chunk = (T + n_splits - 1) / n_splits;
  chunk = (chunk + 31) & ~31;
  const int t_begin = split * chunk;
  const int t_end = min(T, t_begin + chunk);

  float qf[G][8], o[G][8], m[G], l[G];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const uint4 a = *reinterpret_cast<const uint4*>(q + ((long long)b * H + kvh * G + g) * HD + dl);
    qf[g][0] = bf16_lo(a.x) * scale_log2; qf[g][1] = bf16_hi(a.x) * scale_log2;
    qf[g][2] = bf16_lo(a.y) * scale_log2; qf[g][3] = bf16_hi(a.y) * scale_log2;
    qf[g][4] = bf16_lo(a.z) * scale_log2; qf[g][5] = bf16_hi(a.z) * scale_log2;
    qf[g][6] = bf16_lo(a.w) * scale_log2; qf[g][7] = bf16_hi(a.w) * scale_log2;
    m[g] = -INFINITY;
    l[g] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) o[g][j] = 0.f;
  }

  const int32_t* bt = kv.block_table + (long long)b * kv.max_pages;
  // K/V rows of tokens t0 + grp + u * NG, u < UNROLL (zeros past the end of the split)
  auto fetch = [&](int t0, uint4 (&kk)[UNROLL], uint4 (&vv)[UNROLL]) {
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int tt = t0 + grp + u * NG;
      kk[u] = make_uint4(0, 0, 0, 0);
      vv[u] = make_uint4(0, 0, 0, 0);
      if (tt < t_end) {
        const int page = bt[tt / kv.page_tokens], slot = tt % kv.page_tokens;
        const long long off = (((long long)page * KV + kvh) * kv.page_tokens + slot) * HD + dl;
        kk[u] = ld_nc_v4(kv.k + off);
        vv[u] = ld_nc_v4(kv.v + off);
      }
    }
  };
  uint4 kk[UNROLL], vv[UNROLL], nk[UNROLL], nv[UNROLL];
  fetch(t_begin, kk, vv);
  // warp-uniform trip count: every lane takes part in the shuffles of every iteration.  The next iteration's
  // rows are requested before this iteration's arithmetic (register double buffering) so that HBM latency is
  // overlapped instead of paid once per iteration.
  for (int t0 = t_begin; t0 < t_end; t0 += NG * UNROLL) {
    fetch(t0 + NG * UNROLL, nk, nv);
    bool ok[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int tt = t0 + grp + u * NG;
      ok[u] = tt < t_end && (!mask.key_valid || mask.key_valid[(long long)b * mask.key_stride + tt] != 0);  // padded keys
    }
    float s[UNROLL][G];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const float kf[8] = {bf16_lo(kk[u].x), bf16_hi(kk[u].x), bf16_lo(kk[u].y), bf16_hi(kk[u].y),
                           bf16_lo(kk[u].z), bf16_hi(kk[u].z), bf16_lo(kk[u].w), bf16_hi(kk[u].w)};
#pragma unroll
      for (int g = 0; g < G; ++g) {
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) acc = fmaf(qf[g][j], kf[j], acc);
#pragma unroll
        for (int off = LPT / 2; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
        s[u][g] = ok[u] ? acc : -INFINITY;
      }
    }
#pragma unroll
    for (int g = 0; g < G; ++g) {
      float mn = m[g];
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) mn = fmaxf(mn, s[u][g]);
      const float mref = mn == -INFINITY ? 0.f : mn;  // group has seen no token yet: keep everything at zero
      const float corr = exp2f(m[g] - mref);
      m[g] = mn;
      l[g] *= corr;
#pragma unroll
      for (int j = 0; j < 8; ++j) o[g][j] *= corr;
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const float pw = exp2f(s[u][g] - mref);
        l[g] += pw;
        const float vf[8] = {bf16_lo(vv[u].x), bf16_hi(vv[u].x), bf16_lo(vv[u].y), bf16_hi(vv[u].y),
                             bf16_lo(vv[u].z), bf16_hi(vv[u].z), bf16_lo(vv[u].w), bf16_hi(vv[u].w)};
#pragma unroll
        for (int j = 0; j < 8; ++j) o[g][j] = fmaf(pw, vf[j], o[g][j]);
      }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) { kk[u] = nk[u]; vv[u] = nv[u]; }
  }

  // merge the NG token groups of the CTA through shared memory
  __shared__ float sm_o[NG][G][HD];
  __shared__ float sm_m[NG][G], sm_l[NG][G];
#pragma unroll
  for (int g = 0; g < G; ++g) {
#pragma unroll
    for (int j = 0; j < 8; ++j) sm_o[grp][g][dl + j] = o[g][j];
    if ((lane % LPT) == 0) { sm_m[grp][g] = m[g]; sm_l[grp][g] = l[g]; }
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < G * HD; idx += DEC_THREADS) {
    const int g = idx / HD, d = idx - g * HD;
    float M = -INFINITY;
#pragma unroll
    for (int i = 0; i < NG; ++i) M = fmaxf(M, sm_m[i][g]);
    float L = 0.f, O = 0.f;
    if (M > -INFINITY) {
#pragma unroll
      for (int i = 0; i < NG; ++i) {
        const float w = exp2f(sm_m[i][g] - M);  // groups that saw no token have m = -inf -> w = 0
        L = fmaf(sm_l[i][g], w, L);
        O = fmaf(sm_o[i][g][d], w, O);
      }
    }
    if (n_splits == 1) {  // the whole context in one CTA: normalise and store, no combine pass
      out[((long long)b * H + kvh * G + g) * HD + d] = __float2bfloat16_rn(L > 0.f ? O / L : 0.f);
    } else {
      float* dst = scratch + (((long long)b * H + kvh * G + g) * n_splits + split) * (HD + 2);
      dst[d] = O;
      if (d == 0) { dst[HD] = M; dst[HD + 1] = L; }
    }
  }
}


// -------------------------------------------------------------------------------------------------------
// Staged variant (the default).  The direct kernel above keeps at most one 16 KB fetch per CTA in flight
// and every fetch waits on a block-table lookup first, which left HBM half idle (ncu: 50 % DRAM
// throughput, 21 % warps active at B=16, ctx 512).  Here a fifth warp is a pure producer: it walks the
// block table and streams K and V into a ring of DEC_STAGES x 32-token stages with 1-D bulk copies
// (cp.async.bulk, one per page piece, completion counted on the stage's mbarrier), so that 48 KB per CTA
// -- about 24 MB over a full grid -- is in flight whatever the consumer warps are doing.  The four
// consumer warps use the direct kernel's lane mapping and arithmetic; they copy a stage from shared
// memory to registers, hand the slot back at once and only then do the math.
// -------------------------------------------------------------------------------------------------------
constexpr int DEC_STAGES = 3;
constexpr int DEC_STAGE_TOK = 32;
constexpr int DEC_THREADS_STAGED = DEC_THREADS + 32;

template <int HD>
__host__ __device__ constexpr int dec_stage_bytes() { return DEC_STAGE_TOK * HD * 2 * 2; }  // K then V, bf16
// the ring and its barriers, or the end-of-kernel merge buffer that overlays them, whichever is larger
template <int HD, int G>
__host__ __device__ constexpr int dec_smem_bytes() {
  constexpr int ring = DEC_STAGES * dec_stage_bytes<HD>() + 2 * DEC_STAGES * 8;
  constexpr int merge = (DEC_THREADS / 32) * (32 / (HD / 8)) * G * (HD + 2) * 4;
  return ring > merge ? ring : merge;
}

__device__ __forceinline__ void bulk_load_1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ uint4 lds_v4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}

template <int HD, int G>
__global__ void __launch_bounds__(DEC_THREADS_STAGED) decode_attn_kernel(const __nv_bfloat16* __restrict__ q, KvView kv,
                                                                         float* __restrict__ scratch,
                                                                         __nv_bfloat16* __restrict__ out, int H, int KV,
                                                                         float scale_log2, int early_kv,
                                                                         const DecRope rp_in) {
  constexpr int LPT = HD / 8;                   // lanes per token
  constexpr int GPW = 32 / LPT;                 // token groups per warp
  constexpr int NG = (DEC_THREADS / 32) * GPW;  // token groups per CTA
  constexpr int UNROLL = DEC_STAGE_TOK / NG;    // tokens of a stage per group
  constexpr int UH = G <= 2 ? UNROLL : (UNROLL >= 2 ? 2 : 1);  // tokens held in registers at a time
  constexpr int ROW = HD * 2;                   // bytes of one token's K (or V) row
  constexpr int KPART = DEC_STAGE_TOK * ROW;
  constexpr int STAGE = dec_stage_bytes<HD>();
  static_assert(UNROLL % UH == 0, "stage must split evenly");
  static_assert(NG * G * (HD + 2) * 4 <= dec_smem_bytes<HD, G>(), "merge buffer must fit");

  extern __shared__ __align__(128) unsigned char dsm[];
  const uint32_t ring = smem_u32(dsm);
  const uint32_t full0 = ring + DEC_STAGES * STAGE, empty0 = full0 + 8 * DEC_STAGES;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  if (threadIdx.x == 0) {
    for (int i = 0; i < DEC_STAGES; ++i) {
      mbar_init(full0 + 8 * i, 1);
      mbar_init(empty0 + 8 * i, DEC_THREADS / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  pdl_trigger();
  if (threadIdx.x == 0) DLI_TRACE_EV(TR_ATTN, 0);
  // With early_kv the producer warp starts streaming before the dependency wait: sequence lengths, the block table
  // and the pages of every token but this step's own (position T - 1, which the preceding RoPE/append kernel is
  // writing) date from before the current chain of launches.  Everyone else waits here.
  const bool producer = warp == DEC_THREADS / 32;
  if (!(producer && early_kv)) pdl_wait();
  if (threadIdx.x == 0) DLI_TRACE_EV(TR_ATTN, 1);

  const int split = blockIdx.x, n_splits = gridDim.x, kvh = blockIdx.y, b = blockIdx.z;
  const int T = kv.seq_lens[b] + 1;  // the step's own token is (being) appended at position T - 1
  int chunk = (T + n_splits - 1) / n_splits;
  chunk = (chunk + DEC_STAGE_TOK - 1) & ~(DEC_STAGE_TOK - 1);
  const int t_begin = split * chunk;
  const int t_end = min(T, t_begin + chunk);
  const int n_stages = t_end > t_begin ? (t_end - t_begin + DEC_STAGE_TOK - 1) / DEC_STAGE_TOK : 0;
  // fused RoPE + append: the pages hold tokens [0, T - 1); the step's own key and value come from the projection
  // buffer, go to the page AND straight into their row of the last ring stage
  const DecRope rp = at_step_row<HD>(rp_in);
  const bool fused = rp.qkv != nullptr;
  // split context: every CTA rotates its own copy of the query heads; the ONE split whose range holds position T - 1
  // rotates and appends the step's key and value (no other split reads that page slot); the combine kernel re-zeroes
  // the projection buffer once every split has read it
  const bool owns_new = fused && T - 1 >= t_begin && T - 1 < t_end;
  const int t_copy = owns_new ? T - 1 : T;  // tokens the producer takes from the pages

  float o[G][8], m[G], l[G];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    m[g] = -INFINITY;
    l[g] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) o[g][j] = 0.f;
  }
  const int grp = warp * GPW + lane / LPT;  // token group of this lane (consumer warps)
  const int dl = (lane % LPT) * 8;          // head_dim slice [dl, dl + 8)

  if (warp == DEC_THREADS / 32) {
    // ===== producer warp: lane i copies the i-th page piece of the stage =====
    const int32_t* bt = kv.block_table + (long long)b * kv.max_pages;
    const int pt = kv.page_tokens;
    bool waited = !early_kv;
    for (int i = 0; i < n_stages; ++i) {
      const int slot = i % DEC_STAGES;
      const int s0 = t_begin + i * DEC_STAGE_TOK, s1 = min(min(t_end, t_copy), s0 + DEC_STAGE_TOK);
      const int p0 = s0 / pt, p1 = s1 > s0 ? (s1 - 1) / pt : p0 - 1;  // s1 <= s0: the stage holds only the new token
      int page = -1;
      if (p0 + lane <= p1) page = bt[p0 + lane];
      if (!waited && s1 >= T) {  // this stage holds the step's own token: not before the append kernel is done
        pdl_wait();
        waited = true;
      }
      if (i >= DEC_STAGES) mbar_wait(empty0 + 8 * slot, ((uint32_t)(i / DEC_STAGES) - 1) & 1);
      if (lane == 0) mbar_expect_tx(full0 + 8 * slot, (uint32_t)max(s1 - s0, 0) * ROW * 2);
      __syncwarp();
      if (page >= 0) {
        const int ta = max(s0, (p0 + lane) * pt), tb = min(s1, (p0 + lane + 1) * pt);
        const long long off = (((long long)page * KV + kvh) * pt + (ta - (p0 + lane) * pt)) * HD;
        const uint32_t dst = ring + slot * STAGE + (uint32_t)(ta - s0) * ROW;
        const uint32_t bytes = (uint32_t)(tb - ta) * ROW;
        bulk_load_1d(dst, kv.k + off, bytes, full0 + 8 * slot);
        bulk_load_1d(dst + KPART, kv.v + off, bytes, full0 + 8 * slot);
      }
    }
    if (!waited) pdl_wait();  // this warp also stores results below
  } else {
    float qf[G][8];
    uint4 newrow = make_uint4(0, 0, 0, 0);  // fused: this lane's chunk of the step's own K (lanes < LPT) / V row
#pragma unroll
    for (int g = 0; g < G; ++g) {
      uint4 a;
      if (fused) {  // rotate, then round to bf16 exactly as the stand-alone RoPE kernel's q buffer would hold it
        float r8[8];
        rope8<HD>(rp.qkv + (long long)b * rp.n_all + (kvh * G + g) * HD, rp.cos, rp.sin, dl, r8);
        a = make_uint4(pack_bf16(r8[0], r8[1]), pack_bf16(r8[2], r8[3]), pack_bf16(r8[4], r8[5]), pack_bf16(r8[6], r8[7]));
      } else {
        a = *reinterpret_cast<const uint4*>(q + ((long long)b * H + kvh * G + g) * HD + dl);
      }
      qf[g][0] = bf16_lo(a.x) * scale_log2; qf[g][1] = bf16_hi(a.x) * scale_log2;
      qf[g][2] = bf16_lo(a.y) * scale_log2; qf[g][3] = bf16_hi(a.y) * scale_log2;
      qf[g][4] = bf16_lo(a.z) * scale_log2; qf[g][5] = bf16_hi(a.z) * scale_log2;
      qf[g][6] = bf16_lo(a.w) * scale_log2; qf[g][7] = bf16_hi(a.w) * scale_log2;
    }
    const int new_row = (T - 1 - t_begin) % DEC_STAGE_TOK;  // row of the step's token inside the last stage
    if (owns_new) {
      newrow = new_kv_chunk<HD>(rp, b, H, KV, kvh, lane);
      if (warp == 0) append_new_kv<HD>(kv, b, KV, kvh, T - 1, lane, newrow);
    }
    for (int i = 0; i < n_stages; ++i) {
      const int slot = i % DEC_STAGES;
      const int t0 = t_begin + i * DEC_STAGE_TOK;
      const uint32_t kbase = ring + slot * STAGE + (uint32_t)grp * ROW + dl * 2;
      mbar_wait(full0 + 8 * slot, (uint32_t)(i / DEC_STAGES) & 1);
      if (owns_new && i == n_stages - 1 && warp == (new_row % NG) / GPW) {
        // the warp whose token group reads that row puts the new K and V there (the bulk copies never touch it)
        if (lane < 2 * LPT) {
          const uint32_t dst = ring + slot * STAGE + (lane < LPT ? 0u : (uint32_t)KPART) + (uint32_t)new_row * ROW +
                               (uint32_t)(lane % LPT) * 16;
          asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(newrow.x), "r"(newrow.y), "r"(newrow.z),
                       "r"(newrow.w)
                       : "memory");
        }
        __syncwarp();
      }
#pragma unroll
      for (int h = 0; h < UNROLL / UH; ++h) {
        uint4 kk[UH], vv[UH];
        bool ok[UH];
#pragma unroll
        for (int u = 0; u < UH; ++u) {
          const int tok = (h * UH + u) * NG;
          ok[u] = t0 + grp + tok < t_end;  // rows past the end of the split were never copied
          kk[u] = lds_v4(kbase + tok * ROW);
          vv[u] = lds_v4(kbase + KPART + tok * ROW);
          if (!ok[u]) vv[u] = make_uint4(0, 0, 0, 0);
        }
        if (h == UNROLL / UH - 1) {  // the stage is in registers: give the slot back before the arithmetic
          __syncwarp();
          if (lane == 0) mbar_arrive(empty0 + 8 * slot);
        }
        float s[UH][G];
#pragma unroll
        for (int u = 0; u < UH; ++u) {
          const float kf[8] = {bf16_lo(kk[u].x), bf16_hi(kk[u].x), bf16_lo(kk[u].y), bf16_hi(kk[u].y),
                               bf16_lo(kk[u].z), bf16_hi(kk[u].z), bf16_lo(kk[u].w), bf16_hi(kk[u].w)};
#pragma unroll
          for (int g = 0; g < G; ++g) {
            float acc = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) acc = fmaf(qf[g][j], kf[j], acc);
#pragma unroll
            for (int off = LPT / 2; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
            s[u][g] = ok[u] ? acc : -INFINITY;
          }
        }
#pragma unroll
        for (int g = 0; g < G; ++g) {
          float mn = m[g];
#pragma unroll
          for (int u = 0; u < UH; ++u) mn = fmaxf(mn, s[u][g]);
          const float mref = mn == -INFINITY ? 0.f : mn;  // group has seen no token yet: keep everything at zero
          const float corr = exp2f(m[g] - mref);
          m[g] = mn;
          l[g] *= corr;
#pragma unroll
          for (int j = 0; j < 8; ++j) o[g][j] *= corr;
#pragma unroll
          for (int u = 0; u < UH; ++u) {
            const float pw = exp2f(s[u][g] - mref);
            l[g] += pw;
            const float vf[8] = {bf16_lo(vv[u].x), bf16_hi(vv[u].x), bf16_lo(vv[u].y), bf16_hi(vv[u].y),
                                 bf16_lo(vv[u].z), bf16_hi(vv[u].z), bf16_lo(vv[u].w), bf16_hi(vv[u].w)};
#pragma unroll
            for (int j = 0; j < 8; ++j) o[g][j] = fmaf(pw, vf[j], o[g][j]);
          }
        }
      }
    }
  }

  // merge the NG token groups of the CTA through the (now idle) ring
  __syncthreads();
  float* sm_o = reinterpret_cast<float*>(dsm);            // [NG][G][HD]
  float* sm_m = sm_o + NG * G * HD;                       // [NG][G]
  float* sm_l = sm_m + NG * G;                            // [NG][G]
  if (warp < DEC_THREADS / 32) {
#pragma unroll
    for (int g = 0; g < G; ++g) {
#pragma unroll
      for (int j = 0; j < 8; ++j) sm_o[(grp * G + g) * HD + dl + j] = o[g][j];
      if ((lane % LPT) == 0) { sm_m[grp * G + g] = m[g]; sm_l[grp * G + g] = l[g]; }
    }
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < G * HD; idx += DEC_THREADS_STAGED) {
    const int g = idx / HD, d = idx - g * HD;
    float M = -INFINITY;
#pragma unroll
    for (int i = 0; i < NG; ++i) M = fmaxf(M, sm_m[i * G + g]);
    float L = 0.f, O = 0.f;
    if (M > -INFINITY) {
#pragma unroll
      for (int i = 0; i < NG; ++i) {
        const float w = exp2f(sm_m[i * G + g] - M);  // groups that saw no token have m = -inf -> w = 0
        L = fmaf(sm_l[i * G + g], w, L);
        O = fmaf(sm_o[(i * G + g) * HD + d], w, O);
      }
    }
    if (n_splits == 1) {  // the whole context in one CTA: normalise and store, no combine pass
      out[((long long)b * H + kvh * G + g) * HD + d] = __float2bfloat16_rn(O / L);
    } else {
      float* dst = scratch + (((long long)b * H + kvh * G + g) * n_splits + split) * (HD + 2);
      dst[d] = O;
      if (d == 0) { dst[HD] = M; dst[HD + 1] = L; }
    }
  }
  if (fused && n_splits == 1) zero_consumed<HD, G, DEC_THREADS_STAGED>(rp, b, H, KV, kvh);  // every read of the row is behind the barriers above
  if (threadIdx.x == 0) DLI_TRACE_EV(TR_ATTN, 2);
}

template <int HD>
__global__ void __launch_bounds__(HD) decode_combine_kernel(const float* __restrict__ scratch, __nv_bfloat16* __restrict__ out,
                                                            int H, int n_splits, float* __restrict__ qkv_zero, int n_all,
                                                            int KV) {
  pdl_trigger();
  pdl_wait();
  const int h = blockIdx.x, b = blockIdx.y, d = threadIdx.x;
  if (qkv_zero) {  // fused RoPE over a split context: every split has read the projection row by now -> re-zero it
    float* row = qkv_zero + (long long)b * n_all;
    row[h * HD + d] = 0.f;
    const int G = H / KV;
    if (h % G == 0) {
      row[(H + h / G) * HD + d] = 0.f;
      row[(H + KV + h / G) * HD + d] = 0.f;
    }
  }
  const float* src = scratch + ((long long)b * H + h) * n_splits * (HD + 2);
  float M = -INFINITY;
  for (int s = 0; s < n_splits; ++s) M = fmaxf(M, src[s * (HD + 2) + HD]);
  float L = 0.f, O = 0.f;
  for (int s = 0; s < n_splits; ++s) {
    const float w = exp2f(src[s * (HD + 2) + HD] - M);
    L = fmaf(src[s * (HD + 2) + HD + 1], w, L);
    O = fmaf(src[s * (HD + 2) + d], w, O);
  }
  out[((long long)b * H + h) * HD + d] = __float2bfloat16_rn(L > 0.f ? O / L : 0.f);
}

template <int HD, int G>
static int launch_dec(const __nv_bfloat16* q, __nv_bfloat16* out, const KvView& kv, int batch, int H, int KV, int splits,
                      float* scratch, bool early_kv, const DecRope& rope, cudaStream_t st) {
  const float scale_log2 = 1.4426950408889634f / sqrtf((float)HD);
  static const bool direct = exp_env("DLI_DEC_DIRECT") != nullptr;  // A/B switch for the register-staged kernel
  int rc;
  if (direct) {
    rc = launch_kernel("decode_attn_direct_kernel", decode_attn_direct_kernel<HD, G>, dim3(splits, KV, batch),
                       dim3(DEC_THREADS), 0, st, q, kv, scratch, out, H, KV, scale_log2, AttnMask{nullptr, nullptr, 0});
  } else {
    DLI_ONCE_PER_DEVICE(cudaFuncSetAttribute(decode_attn_kernel<HD, G>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       dec_smem_bytes<HD, G>()));  // > 48 KB dynamic shared memory: opt-in per device
    rc = launch_kernel("decode_attn_kernel", decode_attn_kernel<HD, G>, dim3(splits, KV, batch),
                       dim3(DEC_THREADS_STAGED), dec_smem_bytes<HD, G>(), st, q, kv, scratch, out, H, KV, scale_log2,
                       early_kv ? 1 : 0, rope);
  }
  if (rc || splits == 1) return rc;
  return launch_kernel("decode_combine_kernel", decode_combine_kernel<HD>, dim3(H, batch), dim3(HD), 0, st,
                       (const float*)scratch, out, H, splits, const_cast<float*>(rope.qkv), rope.n_all, KV);
}

// The attention kernel can do the RoPE kernel's work when one CTA per (sequence, KV head) is the only reader of that
// head's projection and the only writer of its page slot (unsplit context), or when the context is split: every split
// rotates its own copy of the queries, the split that holds the step's position appends the new key / value, and the
// combine kernel re-zeroes the projection buffer.
bool decode_attention_fuses_rope(const KvView& kv, int batch, int H, int KV, int hd) {
  static const bool off = exp_env("DLI_DEC_DIRECT") != nullptr || exp_env("DLI_NO_ROPE_FUSE") != nullptr;
  if (off || KV <= 0 || H % KV != 0 || !(hd == 64 || hd == 128)) return false;
  const int G = H / KV;
  if (!(G == 1 || G == 2 || G == 4 || G == 8)) return false;
  (void)kv; (void)batch;
  return true;
}

int launch_decode_attention(const __nv_bfloat16* q, __nv_bfloat16* out, const KvView& kv, int batch, int H, int KV, int hd,
                            float* scratch, size_t scratch_bytes, cudaStream_t st, bool early_kv, const DecRope* rope_in) {
  DLI_REQUIRE((q || rope_in) && out && scratch, "dli_decode_attention: null pointer");
  DLI_REQUIRE(batch > 0 && H > 0 && KV > 0 && H % KV == 0, "dli_decode_attention: batch=%d H=%d KV=%d", batch, H, KV);
  const int G = H / KV;
  const bool gqa = decode_attn_gqa_supported(kv, G, hd);
  const int splits = decode_splits_for(kv, batch, H, KV, hd);
  DecRope rope{nullptr, nullptr, nullptr, 0};
  if (rope_in) {
    DLI_REQUIRE(rope_in->qkv && rope_in->cos && rope_in->sin, "decode attention with fused RoPE: null projection / cos / sin");
    rope = *rope_in;
  }
  const size_t need = (size_t)batch * H * splits * (hd + 2) * sizeof(float);
  if (scratch_bytes < need)
    return fail(DLI_ERR_WORKSPACE, "dli_decode_attention: scratch %zu bytes < %zu needed", scratch_bytes, need);
  if (gqa) {  // grouped-query models: tensor-core kernel (attn_decode_gqa.cu)
    int rc = launch_decode_attention_gqa(q, out, kv, batch, H, KV, hd, splits, scratch, early_kv, rope, st);
    if (rc || splits == 1) return rc;
    if (hd == 128)
      return launch_kernel("decode_combine_kernel", decode_combine_kernel<128>, dim3(H, batch), dim3(128), 0, st,
                           (const float*)scratch, out, H, splits, const_cast<float*>(rope.qkv), rope.n_all, KV);
    return launch_kernel("decode_combine_kernel", decode_combine_kernel<64>, dim3(H, batch), dim3(64), 0, st,
                         (const float*)scratch, out, H, splits, const_cast<float*>(rope.qkv), rope.n_all, KV);
  }
#define DLI_DEC_CASE(HD_, G_) \
  if (hd == HD_ && G == G_) return launch_dec<HD_, G_>(q, out, kv, batch, H, KV, splits, scratch, early_kv, rope, st)
  DLI_DEC_CASE(128, 1); DLI_DEC_CASE(128, 2); DLI_DEC_CASE(128, 4); DLI_DEC_CASE(128, 8);
  DLI_DEC_CASE(64, 1); DLI_DEC_CASE(64, 2); DLI_DEC_CASE(64, 4); DLI_DEC_CASE(64, 8);
#undef DLI_DEC_CASE
  return fail(DLI_ERR_UNSUPPORTED, "dli_decode_attention: head_dim %d with GQA group %d is not instantiated", hd, G);
}

// Decode attention with padded keys (dli_decode_attention_masked): the register-staged kernel with a validity check per
// key; a padded batch is the rare path, the tensor-core GQA and bulk-copy kernels stay mask-free.
template <int HD, int G>
static int launch_dec_masked(const __nv_bfloat16* q, __nv_bfloat16* out, const KvView& kv, int batch, int H, int KV, int splits,
                             float* scratch, const AttnMask& mask, cudaStream_t st) {
  const float scale_log2 = 1.4426950408889634f / sqrtf((float)HD);
  int rc = launch_kernel("decode_attn_direct_kernel<masked>", decode_attn_direct_kernel<HD, G>, dim3(splits, KV, batch),
                         dim3(DEC_THREADS), 0, st, q, kv, scratch, out, H, KV, scale_log2, mask);
  if (rc || splits == 1) return rc;
  return launch_kernel("decode_combine_kernel", decode_combine_kernel<HD>, dim3(H, batch), dim3(HD), 0, st,
                       (const float*)scratch, out, H, splits, (float*)nullptr, 0, KV);
}

int launch_decode_attention_masked(const __nv_bfloat16* q, __nv_bfloat16* out, const KvView& kv, int batch, int H, int KV,
                                   int hd, const AttnMask& mask, float* scratch, size_t scratch_bytes, cudaStream_t st) {
  DLI_REQUIRE(q && out && scratch, "dli_decode_attention_masked: null pointer");
  DLI_REQUIRE(batch > 0 && H > 0 && KV > 0 && H % KV == 0, "dli_decode_attention_masked: batch=%d H=%d KV=%d", batch, H, KV);
  DLI_REQUIRE(!mask.key_valid || mask.key_stride >= kv.max_ctx, "dli_decode_attention_masked: key_valid rows of %d entries < context %d",
              mask.key_stride, kv.max_ctx);
  const int G = H / KV;
  const int splits = decode_splits(batch, KV, kv.max_ctx, 2 * sm_count());
  const size_t need = (size_t)batch * H * splits * (hd + 2) * sizeof(float);
  if (scratch_bytes < need)
    return fail(DLI_ERR_WORKSPACE, "dli_decode_attention_masked: scratch %zu bytes < %zu needed", scratch_bytes, need);
#define DLI_DEC_CASE(HD_, G_) \
  if (hd == HD_ && G == G_) return launch_dec_masked<HD_, G_>(q, out, kv, batch, H, KV, splits, scratch, mask, st)
  DLI_DEC_CASE(128, 1); DLI_DEC_CASE(128, 2); DLI_DEC_CASE(128, 4); DLI_DEC_CASE(128, 8);
  DLI_DEC_CASE(64, 1); DLI_DEC_CASE(64, 2); DLI_DEC_CASE(64, 4); DLI_DEC_CASE(64, 8);
#undef DLI_DEC_CASE
  return fail(DLI_ERR_UNSUPPORTED, "dli_decode_attention_masked: head_dim %d with GQA group %d is not instantiated", hd, G);
}

// -------------------------------------------------------------------------------------------------------
// Allocation-free decode attention.
//
// Some users of this library are handed input and output buffers and nothing else: an operator plug-in
// whose host framework has already claimed all of HBM for its own pool, or a serving loop that must not
// hit cudaMalloc's implicit device synchronisation (or its failure when the pool is full) in the middle
// of a request.  They cannot provide the split-K scratch that dli_decode_attention asks for.  For them
// the partials live in the fixed array below, which is part of the library image: every GPU that loads
// libdli_b200.so sets aside DLI_STATIC_SCRATCH_BYTES for it until the process exits.
//
// Host code must not pass `g_decode_scratch` itself to a kernel: on the host that name is only a handle
// for the symbol, not a device address.  The address is looked up once per device with
// cudaGetSymbolAddress while that device is current, and remembered.
// -------------------------------------------------------------------------------------------------------
constexpr size_t DLI_STATIC_SCRATCH_BYTES = (size_t)8 << 20;  // 8 MiB
static __device__ __align__(256) float g_decode_scratch[DLI_STATIC_SCRATCH_BYTES / sizeof(float)];

static float* g_decode_scratch_addr[DLI_MAX_DEVICES] = {nullptr};

static int static_scratch_for_current_device(float** out) {
  int dev = -1;
  DLI_CUDA_OK(cudaGetDevice(&dev));
  DLI_REQUIRE(dev >= 0 && dev < DLI_MAX_DEVICES, "dli_decode_attention_static: device ordinal %d not supported", dev);
  if (g_decode_scratch_addr[dev] == nullptr) {
    void* p = nullptr;
    DLI_CUDA_OK(cudaGetSymbolAddress(&p, g_decode_scratch));
    DLI_REQUIRE(p != nullptr, "dli_decode_attention_static: cudaGetSymbolAddress returned a null address");
    g_decode_scratch_addr[dev] = static_cast<float*>(p);
  }
  *out = g_decode_scratch_addr[dev];
  return DLI_OK;
}

DLI_TRACE_SETTER(trace_set_attn)

}  // namespace dli

extern "C" int dli_decode_attention_static(const void* q_bf16, void* out_bf16, const dli_kv_desc* kv, int layer, int batch,
                                           int n_heads, int n_kv_heads, int head_dim, dli_stream_t stream) {
  using namespace dli;
  DLI_REQUIRE(batch > 0 && n_heads > 0 && n_kv_heads > 0 && head_dim > 0, "dli_decode_attention_static: bad shape");
  KvView v;
  int rc = kv_view(kv, layer, &v);
  if (rc) return rc;
  // same split count the launcher will choose, so the size check here is exact
  const size_t need = (size_t)batch * n_heads * decode_splits_for(v, batch, n_heads, n_kv_heads, head_dim) * (head_dim + 2) * sizeof(float);
  if (need > DLI_STATIC_SCRATCH_BYTES)
    return fail(DLI_ERR_WORKSPACE,
                "dli_decode_attention_static: this problem needs %zu bytes of split-K scratch but the built-in array holds "
                "%zu; use dli_decode_attention with caller-provided scratch",
                need, (size_t)DLI_STATIC_SCRATCH_BYTES);
  float* scratch = nullptr;
  if ((rc = static_scratch_for_current_device(&scratch))) return rc;
  // Nothing has to be cleared first: every partial slot the combine step reads is written by this same call
  // (a split with no tokens still writes m = -inf, l = 0, o = 0), so what the previous call left is never read.
  return launch_decode_attention((const __nv_bfloat16*)q_bf16, (__nv_bfloat16*)out_bf16, v, batch, n_heads, n_kv_heads,
                                 head_dim, scratch, DLI_STATIC_SCRATCH_BYTES, (cudaStream_t)stream);
}
