// tcgen05 flash attention for prefill, second generation (head_dim 128, chunks of >= 256 queries).
//
// Same contractions as attn_prefill_tc.cu (S = Q K^T and O += P V on the 5th-generation tensor cores, fp32 softmax
// statistics in the exp2 domain, lazy rescaling of O in TMEM), re-cut around what limited that kernel: every
// tcgen05.mma with both operands in shared memory occupies the tensor pipe for ~70 cycles whatever its N (operand
// reads), so 64-key tiles paid 12 MMAs per 64 keys and two co-resident CTAs kept the pipe ~80 % occupied at 23 % math
// utilisation.  Here:
//   * key tiles are 128 wide: Q K^T is 8 MMAs of N = 128 per 128 keys (half the issue slots per key);
//   * P never goes through shared memory: the softmax warps write it (bf16 pairs) over the S columns in TMEM and
//     P V is a TS-mode MMA (A operand from TMEM), which also removes the proxy fence and the 32 KB P tile;
//   * one CTA per SM works on TWO query tiles (A, B: 2 x 128 queries of one head) that share every K/V tile: while
//     the 128 softmax threads of one tile are busy, the MMA warp runs the other tile's P V and next Q K^T.
// TMEM (512 columns): S_A/P_A [0,128), O_A [128,256), S_B/P_B [256,384), O_B [384,512).  MMAs of one thread execute in
// issue order, and a tcgen05.commit covers everything issued before it, so "S_X(j) is complete" also means
// "P V_X(j-1) is complete": the softmax may rescale O_X and overwrite P_X(j-1) as soon as s_full fires.
// Warps: 0 = TMA, 1 = MMA issue, 2-9 = softmax of tile A, 10-17 = softmax of tile B.  A query row is shared by TWO
// threads (warps w and w + 4 reach the same TMEM lanes): each owns 64 of the tile's 128 score columns, they exchange
// their row maxima through shared memory once per tile, and each writes its half of P and owns half of O's columns.
// With one thread per row the exp/max/pack of 128 scores (~2300 cycles) sat on the critical path of every tile:
// S_x(j) -> softmax -> P.V_x(j) -> Q.K^T_x(j+1) (measured 1.12x over the first kernel at 8k instead of the ~1.5x the
// MMA count promised).
#include "common.cuh"
#include "tc_ptx.cuh"

#include <math.h>
#include <stdlib.h>

namespace dli {

constexpr int F2_BM = 128;                       // queries per tile (two tiles per CTA)
constexpr int F2_BN = 128;                       // keys per tile
constexpr int F2_HD = 128;
constexpr int F2_THREADS = 576;                 // TMA warp + MMA warp + 2 tiles x 2 column halves x 4 softmax warps
constexpr int F2_Q_BYTES = F2_BM * F2_HD * 2;    // 32 KB per query tile: two 64-column halves of [128 x 128 B]
constexpr int F2_KV_BYTES = F2_BN * F2_HD * 2;   // 32 KB: two halves of [128 keys x 128 B]
constexpr int F2_TILE_SMEM = 2 * F2_Q_BYTES + 4 * F2_KV_BYTES;  // Q_A, Q_B, K[2], V[2] = 192 KB
constexpr int F2_NBARS = 15;
constexpr int F2_XCH_BYTES = 2 * 2 * 2 * F2_BM * 4;  // row-max exchange between the two halves: [parity][tile][half][row]
constexpr int F2_SMEM = F2_TILE_SMEM + 8 * F2_NBARS + 8 + F2_XCH_BYTES + 1024;  // + barriers, TMEM slot, exchange, slack
constexpr float F2_RESCALE_THRESHOLD = 8.0f;     // log2 units

__device__ __forceinline__ uint64_t f2_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__host__ __device__ constexpr uint32_t f2_idesc(int n, bool b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((b_mn_major ? 1u : 0u) << 16) | ((uint32_t)(n >> 3) << 17) |
         ((uint32_t)(128 >> 4) << 24);
}
// D[tmem] (+)= A[tmem] * B[smem]: A = 128 lanes x 16 bf16 (8 columns of bf16 pairs)
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

struct Fa2Params {
  __nv_bfloat16* out;
  const int32_t* block_table;
  const int32_t* seq_lens;
  int max_pages, page_tokens, q_len, H, KV, causal;
  float scale_log2;
  unsigned long long* dbg;  // optional cycle counters of one CTA (experiments only; nullptr in production)
};

__global__ void __launch_bounds__(F2_THREADS, 1)
prefill_attn_tc2_kernel(const __grid_constant__ CUtensorMap tma_q, const __grid_constant__ CUtensorMap tma_k,
                        const __grid_constant__ CUtensorMap tma_v, const Fa2Params p) {
  extern __shared__ uint8_t f2_smem_raw[];
  const uint32_t base = (smem_u32(f2_smem_raw) + 1023u) & ~1023u;  // the 128B swizzle needs 1024-byte aligned tiles
  uint8_t* aligned = f2_smem_raw + (base - smem_u32(f2_smem_raw));
  // barriers: 0 q_full | 1-2 k_full | 3-4 k_empty | 5-6 v_full | 7-8 v_empty | 9-10 s_full[A,B] | 11-12 p_full[A,B] | 13-14 pv_done[A,B]
  uint64_t* bars = reinterpret_cast<uint64_t*>(aligned + F2_TILE_SMEM);
  uint32_t& tmem_base_slot = *reinterpret_cast<uint32_t*>(aligned + F2_TILE_SMEM + 8 * F2_NBARS);

  const uint32_t sQ = base, sK = sQ + 2 * F2_Q_BYTES, sV = sK + 2 * F2_KV_BYTES;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t bar0 = smem_u32(bars);
  const uint32_t q_full = bar0, k_full = bar0 + 8, k_empty = bar0 + 24, v_full = bar0 + 40, v_empty = bar0 + 56,
                 s_full = bar0 + 72, p_full = bar0 + 88, pv_done = bar0 + 104;

  pdl_trigger();
  if (threadIdx.x == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_q) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_k) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_v) : "memory");
    mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(k_full + 8 * i, 1);
      mbar_init(k_empty + 8 * i, 1);
      mbar_init(v_full + 8 * i, 1);
      mbar_init(v_empty + 8 * i, 1);
      mbar_init(s_full + 8 * i, 1);
      mbar_init(p_full + 8 * i, 256);
      mbar_init(pv_done + 8 * i, 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)),
                 "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem = tmem_base_slot;

  pdl_wait();  // q, the pages and seq_lens are produced by the preceding kernels of the stream
  // causal: the long query tiles of each head are launched first (see attn_prefill_tc.cu)
  const int m0 = (p.causal ? (int)(gridDim.x - 1 - blockIdx.x) : (int)blockIdx.x) * (2 * F2_BM), h = blockIdx.y, b = blockIdx.z;
  const int kvh = h / (p.H / p.KV);
  const int past = p.seq_lens[b];
  const int T = past + p.q_len;
  // key tiles each query tile needs (tile B sees 128 more keys than tile A under the causal mask)
  const int n_a = ((p.causal ? min(T, past + m0 + F2_BM) : T) + F2_BN - 1) / F2_BN;
  const int n_b = ((p.causal ? min(T, past + m0 + 2 * F2_BM) : T) + F2_BN - 1) / F2_BN;

  if (warp == 0) {
    // ===== TMA producer: both query tiles, then K(j+1) and V(j) alternately =====
    const int32_t* bt = p.block_table + (long long)b * p.max_pages;
    const int pages_per_tile = F2_BN / p.page_tokens;
    const uint32_t page_bytes = (uint32_t)p.page_tokens * 128u;  // one head-dim half of one page
    auto lookup = [&](int j) {  // lane pg: first pool row of page pg of tile j (past the end: any valid page, masked later)
      int my_row = 0;
      if (lane < pages_per_tile && j < n_b) {
        const int tok0 = (j * pages_per_tile + lane) * p.page_tokens;
        const int page = tok0 < T ? bt[j * pages_per_tile + lane] : bt[0];
        my_row = (page * p.KV + kvh) * p.page_tokens;
      }
      return my_row;
    };
    auto load_tile = [&](int j, int my_row, const CUtensorMap* map, uint32_t smem_base, uint32_t full_bar) {
      const uint32_t dst = smem_base + (j & 1) * F2_KV_BYTES, bar = full_bar + 8 * (j & 1);
      if (elect_one()) mbar_expect_tx(bar, F2_KV_BYTES);
      __syncwarp();
      for (int pg = 0; pg < pages_per_tile; ++pg) {
        const int row = __shfl_sync(0xffffffffu, my_row, pg);
        if (elect_one()) {
          tma_load_2d(dst + pg * page_bytes, map, bar, 0, row);
          tma_load_2d(dst + F2_KV_BYTES / 2 + pg * page_bytes, map, bar, 64, row);
        }
        __syncwarp();
      }
    };
    if (elect_one()) {
      mbar_expect_tx(q_full, 2 * F2_Q_BYTES);
      for (int x = 0; x < 2; ++x) {
        tma_load_2d(sQ + x * F2_Q_BYTES, &tma_q, q_full, h * F2_HD, b * p.q_len + m0 + x * F2_BM);
        tma_load_2d(sQ + x * F2_Q_BYTES + F2_Q_BYTES / 2, &tma_q, q_full, h * F2_HD + 64, b * p.q_len + m0 + x * F2_BM);
      }
    }
    __syncwarp();
    int row_cur = lookup(0), row_next = lookup(1);
    load_tile(0, row_cur, &tma_k, sK, k_full);
    for (int j = 0; j < n_b; ++j) {
      const int row_next2 = lookup(j + 2);
      if (j + 1 < n_b) {  // K(j+1): its buffer is free once both Q.K^T(j-1) have completed
        if (j >= 1) mbar_wait(k_empty + 8 * ((j + 1) & 1), (((j + 1) >> 1) - 1) & 1, 64);
        load_tile(j + 1, row_next, &tma_k, sK, k_full);
      }
      if (j >= 2) mbar_wait(v_empty + 8 * (j & 1), ((j >> 1) - 1) & 1, 64);  // V(j): free once both P.V(j-2) have completed
      load_tile(j, row_cur, &tma_v, sV, v_full);
      row_cur = row_next;
      row_next = row_next2;
    }
  } else if (warp == 1) {
    // ===== MMA issuer: per key tile, for A then B: P.V(j) of the tile, then its Q.K^T(j+1) =====
    constexpr uint32_t idesc_qk = f2_idesc(F2_BN, false);
    constexpr uint32_t idesc_pv = f2_idesc(F2_HD, true);
    auto issue_qk = [&](int x, int j) {  // S_x = Q_x . K(j)^T
      if (elect_one()) {
        const int kbuf = j & 1;
#pragma unroll
        for (int kb = 0; kb < 2; ++kb) {
          const uint64_t a = umma_desc_sw128(sQ + x * F2_Q_BYTES + kb * (F2_Q_BYTES / 2)),
                         bd = umma_desc_sw128(sK + kbuf * F2_KV_BYTES + kb * (F2_KV_BYTES / 2));
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16(tmem + x * 256, a + 2 * k, bd + 2 * k, idesc_qk, (kb | k) ? 1u : 0u);
        }
        umma_commit(s_full + 8 * x);
      }
      __syncwarp();
    };
    auto issue_pv = [&](int x, int j) {  // O_x += P_x . V(j), P from TMEM (8 columns of bf16 pairs per 16 keys)
      if (elect_one()) {
        const int vb = j & 1;
#pragma unroll
        for (int k = 0; k < F2_BN / 16; ++k) {
          const uint64_t bd = f2_desc_mn_sw128(sV + vb * F2_KV_BYTES + k * 2048, F2_KV_BYTES / 2, 1024);
          umma_bf16_ts(tmem + x * 256 + 128, tmem + x * 256 + 8 * k, bd, idesc_pv, (j | k) ? 1u : 0u);
        }
        umma_commit(pv_done + 8 * x);
      }
      __syncwarp();
    };
    const bool mdbg = p.dbg != nullptr && m0 == (int)(gridDim.x - 1) * 2 * F2_BM && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0;
    unsigned long long mc[5] = {0, 0, 0, 0, 0}, mt0 = 0, mt1;
#define F2_MTICK(i) do { if (mdbg) { mt1 = clock64(); mc[i] += mt1 - mt0; mt0 = mt1; } } while (0)
    mbar_wait(q_full, 0, 32);
    mbar_wait(k_full, 0, 32);
    tcgen05_fence_after();
    if (mdbg) mt0 = clock64();
    issue_qk(0, 0);
    issue_qk(1, 0);
    if (elect_one()) umma_commit(k_empty);
    __syncwarp();
    for (int j = 0; j < n_b; ++j) {
      const int vb = j & 1, kn = (j + 1) & 1;
      mbar_wait(v_full + 8 * vb, (j >> 1) & 1, 32);
      F2_MTICK(0);
      if (j < n_a) {
        mbar_wait(p_full, j & 1, 32);
        F2_MTICK(1);
        tcgen05_fence_after();
        issue_pv(0, j);
        if (j + 1 < n_a) {
          mbar_wait(k_full + 8 * kn, ((j + 1) >> 1) & 1, 32);
          F2_MTICK(2);
          tcgen05_fence_after();
          issue_qk(0, j + 1);
        }
        F2_MTICK(4);
      }
      mbar_wait(p_full + 8, j & 1, 32);
      F2_MTICK(3);
      tcgen05_fence_after();
      issue_pv(1, j);
      if (elect_one()) umma_commit(v_empty + 8 * vb);
      __syncwarp();
      if (j + 1 < n_b) {
        mbar_wait(k_full + 8 * kn, ((j + 1) >> 1) & 1, 32);
        tcgen05_fence_after();
        issue_qk(1, j + 1);
        if (elect_one()) umma_commit(k_empty + 8 * kn);
        __syncwarp();
      }
      F2_MTICK(4);
    }
    if (mdbg) {
      for (int i = 0; i < 5; ++i) p.dbg[8 + i] = mc[i];
    }
  } else {
    // ===== softmax warps: tile x = (warp - 2) / 8, column half = ((warp - 2) / 4) & 1, one query row per thread pair =====
    const int x = (warp - 2) >> 3, half = ((warp - 2) >> 2) & 1, q4 = warp & 3;
    const int row = q4 * 32 + lane;
    const int cb = half * (F2_BN / 2);  // first score column of this thread
    const uint32_t lane_addr = (uint32_t)(q4 * 32) << 16;
    const uint32_t tS = tmem + x * 256 + lane_addr, tO = tS + 128;
    const int qi = m0 + x * F2_BM + row;
    const int pos = past + qi;  // absolute position of this query
    const int n_x = x ? n_b : n_a;
    const int row0_pos = past + m0 + x * F2_BM;  // position of the tile's first query
    float* xch = reinterpret_cast<float*>(aligned + F2_TILE_SMEM + 8 * F2_NBARS + 8);  // [parity][tile][half][row]
    float m_ref = -INFINITY, l = 0.f;
    uint32_t v[16];
    const bool dbg_on = p.dbg != nullptr && m0 == (int)(gridDim.x - 1) * 2 * F2_BM && blockIdx.y == 0 && blockIdx.z == 0 &&
                        x == 1 && half == 0 && row == 0;
    unsigned long long tc[5] = {0, 0, 0, 0, 0}, t0 = 0, t1;
#define F2_TICK(i) do { if (dbg_on) { t1 = clock64(); tc[i] += t1 - t0; t0 = t1; } } while (0)
    if (dbg_on) t0 = clock64();
    for (int j = 0; j < n_x; ++j) {
      const int j0 = j * F2_BN + cb;
      float s[F2_BN / 2];  // raw scores (unscaled): 1/sqrt(hd)*log2(e) is folded into the exp2 argument
      mbar_wait(s_full + 8 * x, j & 1, 20);
      if (j > 0) mbar_wait(pv_done + 8 * x, (j - 1) & 1);  // already complete (commit order); keeps the phase in step
      F2_TICK(0);
      tcgen05_fence_after();
#pragma unroll
      for (int c0 = 0; c0 < F2_BN / 2; c0 += 16) {
        tmem_ld16(tS + cb + c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) s[c0 + i] = __uint_as_float(v[i]);
      }
      F2_TICK(1);
      if ((j0 + F2_BN / 2 > T) || (p.causal && j0 + F2_BN / 2 - 1 > row0_pos)) {  // touches the diagonal / context end
#pragma unroll
        for (int i = 0; i < F2_BN / 2; ++i) {
          const int key = j0 + i;
          if (!(key < T && (!p.causal || key <= pos))) s[i] = -INFINITY;
        }
      }
      float mx0 = s[0], mx1 = s[1], mx2 = s[2], mx3 = s[3];
#pragma unroll
      for (int i = 4; i < F2_BN / 2; i += 4) {
        mx0 = fmaxf(mx0, s[i]); mx1 = fmaxf(mx1, s[i + 1]); mx2 = fmaxf(mx2, s[i + 2]); mx3 = fmaxf(mx3, s[i + 3]);
      }
      float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
      // row maximum over both halves.  The barrier also orders "both threads have read their S columns" before either
      // overwrites them with P (the upper half's P lands on columns the lower half reads its scores from).
      float* slot = xch + (((j & 1) * 2 + x) * 2) * F2_BM;
      slot[half * F2_BM + row] = mx;
      F2_TICK(2);
      asm volatile("bar.sync %0, %1;" ::"r"(1 + x), "r"(256) : "memory");
      F2_TICK(3);
      mx = fmaxf(mx, slot[(half ^ 1) * F2_BM + row]) * p.scale_log2;  // scale > 0: max commutes with it
      float corr = 1.f;
      if (mx > m_ref + F2_RESCALE_THRESHOLD || m_ref == -INFINITY) {  // move the reference maximum (same in both threads)
        const float m_new = fmaxf(mx, m_ref);
        if (m_new > -INFINITY) {
          corr = m_ref == -INFINITY ? 1.f : exp2f(m_ref - m_new);
          l *= corr;
          m_ref = m_new;
        }
      }
      if (j > 0 && __any_sync(0xffffffffu, corr != 1.f)) {  // rare: rescale this warp's rows of its half of O in TMEM
#pragma unroll 1
        for (int c0 = 0; c0 < F2_HD / 2; c0 += 16) {
          tmem_ld16(tO + cb + c0, v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) * corr);
          tmem_st16(tO + cb + c0, v);
        }
      }
      const float neg_mref = m_ref == -INFINITY ? 0.f : -m_ref;
      float l0 = 0.f, l1 = 0.f;
#pragma unroll
      for (int c0 = 0; c0 < F2_BN / 2; c0 += 32) {  // 32 scores -> 16 columns of bf16 pairs
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float p0, p1;
          asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p0) : "f"(fmaf(s[c0 + i], p.scale_log2, neg_mref)));
          asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p1) : "f"(fmaf(s[c0 + i + 1], p.scale_log2, neg_mref)));
          l0 += p0;
          l1 += p1;
          v[i / 2] = pack_bf16(p0, p1);
        }
        tmem_st16(tS + (cb + c0) / 2, v);
      }
      l += l0 + l1;
      tmem_st_wait();
      tcgen05_fence_before();
      mbar_arrive(p_full + 8 * x);
      F2_TICK(4);
    }
    if (dbg_on) {
      for (int i = 0; i < 5; ++i) p.dbg[i] = tc[i];
      p.dbg[5] = n_x;
    }

    // ---- finalize: O / l -> bf16 (each thread its half of the head dimension; l is the sum of both halves) ----
    float* slot = xch + (((n_x & 1) * 2 + x) * 2) * F2_BM;
    slot[half * F2_BM + row] = l;
    asm volatile("bar.sync %0, %1;" ::"r"(1 + x), "r"(256) : "memory");
    l += slot[(half ^ 1) * F2_BM + row];
    mbar_wait(pv_done + 8 * x, (n_x - 1) & 1);
    tcgen05_fence_after();
    const float inv = l > 0.f ? 1.f / l : 0.f;
    __nv_bfloat16* dst = p.out + (((long long)b * p.q_len + qi) * p.H + h) * F2_HD + cb;
#pragma unroll 1
    for (int c0 = 0; c0 < F2_HD / 2; c0 += 16) {
      tmem_ld16(tO + cb + c0, v);
      tmem_ld_wait();
      if (qi < p.q_len) {
        uint32_t w[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) w[i] = pack_bf16(__uint_as_float(v[2 * i]) * inv, __uint_as_float(v[2 * i + 1]) * inv);
        *reinterpret_cast<uint4*>(dst + c0) = make_uint4(w[0], w[1], w[2], w[3]);
        *reinterpret_cast<uint4*>(dst + c0 + 8) = make_uint4(w[4], w[5], w[6], w[7]);
      }
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
  }
}

bool prefill_attn_tc2_supported(const KvView& kv, int hd, int q_len) {
  static const bool off = exp_env("DLI_FA_V1") != nullptr || (exp_env("DLI_ATTN_LEGACY") && exp_env("DLI_ATTN_LEGACY")[0] == '1');
  // Long chunks only: with few key tiles per CTA the one-CTA-per-SM kernel pays its fill and drain (16 x 512 tokens:
  // 5.6 ms against 4.6 ms for attn_prefill_tc.cu); at 8k it is ~10 % faster inside the stage (20.0 vs 22.3 ms per 10 layers).
  static const int min_q = exp_env("DLI_FA_V2_MIN_Q") ? atoi(exp_env("DLI_FA_V2_MIN_Q")) : 2048;
  return !off && hd == F2_HD && q_len >= min_q && q_len >= 2 * F2_BM && kv.page_tokens >= 8 && kv.page_tokens % 8 == 0 &&
         F2_BN % kv.page_tokens == 0;
}

int launch_prefill_attention_tc2(const __nv_bfloat16* q, __nv_bfloat16* out, const KvView& kv, int batch, int q_len, int H,
                                 int KV, bool causal, cudaStream_t st) {
  CUtensorMap tq, tk, tv;
  int rc = make_tmap_bf16(&tq, q, (long long)batch * q_len, (long long)H * F2_HD, (long long)H * F2_HD, F2_BM);
  if (rc) return rc;
  const long long pool_rows = (long long)kv.n_pages * KV * kv.page_tokens;
  if ((rc = make_tmap_bf16(&tk, kv.k, pool_rows, F2_HD, F2_HD, kv.page_tokens))) return rc;
  if ((rc = make_tmap_bf16(&tv, kv.v, pool_rows, F2_HD, F2_HD, kv.page_tokens))) return rc;
  DLI_ONCE_PER_DEVICE(cudaFuncSetAttribute(prefill_attn_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, F2_SMEM));  // > 48 KB dynamic shared memory: opt-in per device
  Fa2Params p{};
  p.out = out;
  p.block_table = kv.block_table;
  p.seq_lens = kv.seq_lens;
  p.max_pages = kv.max_pages;
  p.page_tokens = kv.page_tokens;
  p.q_len = q_len;
  p.H = H;
  p.KV = KV;
  p.causal = causal ? 1 : 0;
  p.scale_log2 = 1.4426950408889634f / sqrtf((float)F2_HD);
  {
    const char* e = exp_env("DLI_FA_DBG_PTR");  // experiments: device pointer (hex) receiving cycle counters
    p.dbg = e ? reinterpret_cast<unsigned long long*>(strtoull(e, nullptr, 16)) : nullptr;
  }
  dim3 grid((q_len + 2 * F2_BM - 1) / (2 * F2_BM), H, batch);
  return launch_kernel("prefill_attn_tc2_kernel", prefill_attn_tc2_kernel, grid, dim3(F2_THREADS), (size_t)F2_SMEM, st, tq, tk,
                       tv, p);
}

}  // namespace dli
