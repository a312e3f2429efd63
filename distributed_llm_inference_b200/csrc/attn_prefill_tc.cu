// tcgen05 flash attention for prefill (head_dim 128) over the paged KV cache.
//
// Replaces modules.py:87-105 for multi-token chunks, like attn_prefill.cu, but with both contractions on the
// 5th-generation tensor cores:
//   S = Q K^T   : tcgen05.mma, A = Q tile [128 queries x 128] (K-major, TMA), B = K tile [64 keys x 128] (K-major,
//                 TMA straight from the pages: one box per page and head-dim half), D = S in TMEM (64 fp32 columns)
//   O += P V    : tcgen05.mma, A = P [128 x 64 keys] bf16 written by the softmax warps into 128B-swizzled smem,
//                 B = V tile [64 keys x 128] in its natural [key][head_dim] layout (MN-major descriptor), D = O in TMEM
// Softmax runs on 4 warps, one query row per thread (tcgen05.ld of its S row), fp32 statistics in the exp2 domain.
// O stays in TMEM for the whole key loop; it is rescaled only when a row's running maximum has grown by more than
// 2^8 since its reference maximum was fixed (lazy rescaling), so the TMEM round trip is rare.
// K and V are double-buffered and K runs one tile ahead of V (the order in which their buffers become free).
// Single-thread instructions (TMA, tcgen05.mma, tcgen05.commit) are issued by one ELECTED lane of a CONVERGED warp:
// behind `if (lane == 0)` ptxas wraps each of them in an ELECT/BRA.U.ANY loop that costs ~145 cycles per issue
// (tests/exp/mma_chain.cu), which made the issuing lanes the bottleneck (cycle counters: the Q.K^T issuer waited
// 1459 of 2386 cycles per tile for K; tensor pipe 20 % busy).
// One CTA = 128 queries of one head; 112 KB of shared memory and 256 TMEM columns, so two CTAs share an SM and one
// runs its MMAs while the other is in softmax.  Warp 0 issues TMA, warp 1 issues MMAs, warps 2-5 are the softmax.
#include "common.cuh"
#include "tc_ptx.cuh"

#include <math.h>
#include <stdlib.h>

namespace dli {

constexpr int FA_BM = 128;      // queries per CTA
constexpr int FA_BN = 64;       // keys per tile
constexpr int FA_HD = 128;
constexpr int FA_THREADS = 192;
constexpr int FA_Q_BYTES = FA_BM * FA_HD * 2;        // 32 KB: two 64-column halves of [128 rows x 128 B]
constexpr int FA_KV_BYTES = FA_BN * FA_HD * 2;       // 16 KB: two halves of [64 rows x 128 B]
constexpr int FA_P_BYTES = FA_BM * FA_BN * 2;        // 16 KB: [128 rows x 128 B]
constexpr int FA_TILE_SMEM = FA_Q_BYTES + 4 * FA_KV_BYTES + FA_P_BYTES;  // Q, K[2], V[2], P = 112 KB
constexpr int FA_SMEM = FA_TILE_SMEM + 128;                               // + mbarriers and the TMEM slot
constexpr uint32_t FA_TMEM_COLS = 256;               // S: columns [0, 64), O: columns [64, 192)
constexpr float FA_RESCALE_THRESHOLD = 8.0f;         // log2 units

// MN-major (here: head_dim-contiguous) 128B-swizzled operand: 64-element rows of 128 bytes, 8 rows (keys) per
// 1024-byte swizzle atom; SBO = distance between 8-key groups, LBO = distance between 64-element column blocks.
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// kind::f16 instruction descriptor, D fp32, A/B bf16, A K-major, B K-major or MN-major, M = 128
__host__ __device__ constexpr uint32_t fa_idesc(int n, bool b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((b_mn_major ? 1u : 0u) << 16) | ((uint32_t)(n >> 3) << 17) |
         ((uint32_t)(128 >> 4) << 24);
}

struct FaParams {
  __nv_bfloat16* out;
  const int32_t* block_table;
  const int32_t* seq_lens;
  int max_pages, page_tokens, q_len, H, KV, causal;
  float scale_log2;
  unsigned long long* dbg;  // optional cycle counters of one CTA (experiments only; nullptr in production)
};

__global__ void __launch_bounds__(FA_THREADS, 2)
prefill_attn_tc_kernel(const __grid_constant__ CUtensorMap tma_q, const __grid_constant__ CUtensorMap tma_k,
                       const __grid_constant__ CUtensorMap tma_v, const FaParams p) {
  // No static shared memory: the dynamic window then starts at the CTA's (1024-byte aligned) shared base, and two
  // CTAs of 112 KB + 128 B fit one SM.  Layout: Q | K[2] | V[2] | P | mbarriers | TMEM slot.
  extern __shared__ __align__(1024) uint8_t fa_smem_raw[];
  // barriers: 0 q_full, 1 s_full, 2 s_empty, 3 p_full, 4 pv_done, 5-6 k_full[2], 7-8 k_empty[2], 9-10 v_full[2], 11-12 v_empty[2]
  uint64_t* bars = reinterpret_cast<uint64_t*>(fa_smem_raw + FA_TILE_SMEM);
  uint32_t& tmem_base_slot = *reinterpret_cast<uint32_t*>(fa_smem_raw + FA_TILE_SMEM + 13 * 8);

  const uint32_t base = smem_u32(fa_smem_raw);
  if (base & 1023u) __trap();  // the 128B swizzle needs 1024-byte aligned tiles
  const uint32_t sQ = base, sK = sQ + FA_Q_BYTES, sV = sK + 2 * FA_KV_BYTES, sP = sV + 2 * FA_KV_BYTES;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t q_full = smem_u32(&bars[0]), s_full = smem_u32(&bars[1]), s_empty = smem_u32(&bars[2]),
                 p_full = smem_u32(&bars[3]), pv_done = smem_u32(&bars[4]), k_full = smem_u32(&bars[5]),
                 k_empty = smem_u32(&bars[7]), v_full = smem_u32(&bars[9]), v_empty = smem_u32(&bars[11]);

  pdl_trigger();
  if (threadIdx.x == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_q) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_k) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_v) : "memory");
    mbar_init(q_full, 1);
    mbar_init(k_full, 1);
    mbar_init(k_full + 8, 1);
    mbar_init(k_empty, 1);
    mbar_init(k_empty + 8, 1);
    mbar_init(v_full, 1);
    mbar_init(v_full + 8, 1);
    mbar_init(v_empty, 1);
    mbar_init(v_empty + 8, 1);
    mbar_init(s_full, 1);
    mbar_init(s_empty, 128);
    mbar_init(p_full, 128);
    mbar_init(pv_done, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)),
                 "r"(FA_TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem = tmem_base_slot;
  const uint32_t tS = tmem, tO = tmem + FA_BN;

  pdl_wait();  // q, the pages and seq_lens are produced by the preceding kernels of the stream
  // causal: a later query tile has more key tiles; launch the long ones of each head first so the grid's last wave is
  // made of short CTAs (the q-tile index stays the fastest grid dimension: tiles of one head share its K/V in L2)
  const int m0 = (p.causal ? (int)(gridDim.x - 1 - blockIdx.x) : (int)blockIdx.x) * FA_BM, h = blockIdx.y, b = blockIdx.z;
  const int kvh = h / (p.H / p.KV);
  const int past = p.seq_lens[b];
  const int T = past + p.q_len;
  int kv_end = T;
  if (p.causal) kv_end = min(T, past + m0 + FA_BM);
  const int n_tiles = (kv_end + FA_BN - 1) / FA_BN;

  if (warp == 0) {
    // ===== TMA producer: Q, then K(j+1) and V(j) alternately — the order in which their buffers become free =====
    const int32_t* bt = p.block_table + (long long)b * p.max_pages;
    const int pages_per_tile = FA_BN / p.page_tokens;
    const uint32_t page_bytes = (uint32_t)p.page_tokens * 128u;  // one head-dim half of one page
    auto lookup = [&](int j) {  // lane pg: first pool row of page pg of tile j (past the end: any valid page, masked later)
      int my_row = 0;
      if (lane < pages_per_tile && j < n_tiles) {
        const int tok0 = (j * pages_per_tile + lane) * p.page_tokens;
        const int page = tok0 < T ? bt[j * pages_per_tile + lane] : bt[0];
        my_row = (page * p.KV + kvh) * p.page_tokens;
      }
      return my_row;
    };
    auto load_tile = [&](int j, int my_row, const CUtensorMap* map, uint32_t smem_base, uint32_t full_bar) {
      const uint32_t dst = smem_base + (j & 1) * FA_KV_BYTES, bar = full_bar + 8 * (j & 1);
      if (elect_one()) mbar_expect_tx(bar, FA_KV_BYTES);
      __syncwarp();
      for (int pg = 0; pg < pages_per_tile; ++pg) {
        const int row = __shfl_sync(0xffffffffu, my_row, pg);
        if (elect_one()) {
          tma_load_2d(dst + pg * page_bytes, map, bar, 0, row);
          tma_load_2d(dst + FA_KV_BYTES / 2 + pg * page_bytes, map, bar, 64, row);
        }
        __syncwarp();
      }
    };
    if (elect_one()) {
      mbar_expect_tx(q_full, FA_Q_BYTES);
      tma_load_2d(sQ, &tma_q, q_full, h * FA_HD, b * p.q_len + m0);
      tma_load_2d(sQ + FA_Q_BYTES / 2, &tma_q, q_full, h * FA_HD + 64, b * p.q_len + m0);
    }
    __syncwarp();
    int row_cur = lookup(0), row_next = lookup(1);
    load_tile(0, row_cur, &tma_k, sK, k_full);
    for (int j = 0; j < n_tiles; ++j) {
      const int row_next2 = lookup(j + 2);  // page-table reads stay two tiles ahead of their use
      if (j + 1 < n_tiles) {                // K(j+1): its buffer is free once Q.K^T(j-1) has completed
        if (j >= 1) mbar_wait(k_empty + 8 * ((j + 1) & 1), (((j + 1) >> 1) - 1) & 1, 64);
        load_tile(j + 1, row_next, &tma_k, sK, k_full);
      }
      if (j >= 2) mbar_wait(v_empty + 8 * (j & 1), ((j >> 1) - 1) & 1, 64);  // V(j): free once P.V(j-2) has completed
      load_tile(j, row_cur, &tma_v, sV, v_full);
      row_cur = row_next;
      row_next = row_next2;
    }
  } else if (warp == 1) {
    // ===== MMA issuer (converged warp, elected lane): Q.K^T runs one tile ahead of P.V =====
    constexpr uint32_t idesc_qk = fa_idesc(FA_BN, false);
    constexpr uint32_t idesc_pv = fa_idesc(FA_HD, true);
    auto issue_qk = [&](int j) {
      const int kbuf = j & 1;
      mbar_wait(k_full + 8 * kbuf, (j >> 1) & 1, 32);
      if (j > 0) mbar_wait(s_empty, (j - 1) & 1, 32);  // softmax has copied S of the previous tile to registers
      tcgen05_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int kb = 0; kb < 2; ++kb) {
          const uint64_t a = umma_desc_sw128(sQ + kb * (FA_Q_BYTES / 2)),
                         bd = umma_desc_sw128(sK + kbuf * FA_KV_BYTES + kb * (FA_KV_BYTES / 2));
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16(tS, a + 2 * k, bd + 2 * k, idesc_qk, (kb | k) ? 1u : 0u);
        }
        umma_commit(s_full);
        umma_commit(k_empty + 8 * kbuf);
      }
      __syncwarp();
    };
    mbar_wait(q_full, 0, 32);
    issue_qk(0);
    for (int j = 0; j < n_tiles; ++j) {
      if (j + 1 < n_tiles) issue_qk(j + 1);  // overlaps the softmax of tile j
      const int vb = j & 1;
      mbar_wait(v_full + 8 * vb, (j >> 1) & 1, 32);
      mbar_wait(p_full, j & 1, 32);
      tcgen05_fence_after();
      if (elect_one()) {
        const uint64_t a = umma_desc_sw128(sP);
#pragma unroll
        for (int k = 0; k < 4; ++k) {  // 16 keys per instruction: two 8-key swizzle atoms of V
          const uint64_t bd = umma_desc_mn_sw128(sV + vb * FA_KV_BYTES + k * 2048, FA_KV_BYTES / 2, 1024);
          umma_bf16(tO, a + 2 * k, bd, idesc_pv, (j | k) ? 1u : 0u);
        }
        umma_commit(pv_done);
        umma_commit(v_empty + 8 * vb);
      }
      __syncwarp();
    }
  } else {
    // ===== softmax warps: one query row per thread =====
    const int q4 = warp & 3;
    const int row = q4 * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(q4 * 32) << 16;
    const int qi = m0 + row;
    const int pos = past + qi;  // absolute position of this query
    float m_ref = -INFINITY, l = 0.f;
    uint32_t v[16];
    const bool dbg_on = p.dbg != nullptr && m0 == (int)(gridDim.x - 1) * FA_BM && blockIdx.y == 0 && blockIdx.z == 0 && threadIdx.x == 64;
    unsigned long long tc[5] = {0, 0, 0, 0, 0}, t0 = 0, t1;
#define FA_TICK(i) do { if (dbg_on) { t1 = clock64(); tc[i] += t1 - t0; t0 = t1; } } while (0)
    if (dbg_on) t0 = clock64();
    for (int j = 0; j < n_tiles; ++j) {
      const int j0 = j * FA_BN;
      float s[FA_BN];  // raw scores (unscaled): the 1/sqrt(hd)*log2(e) factor is folded into the exp2 argument
      mbar_wait(s_full, j & 1, 20);
      FA_TICK(0);
      tcgen05_fence_after();
#pragma unroll
      for (int c0 = 0; c0 < FA_BN; c0 += 16) {
        tmem_ld16(tS + lane_addr + c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) s[c0 + i] = __uint_as_float(v[i]);
      }
      tcgen05_fence_before();
      mbar_arrive(s_empty);
      FA_TICK(1);

      if ((j0 + FA_BN > T) || (p.causal && j0 + FA_BN - 1 > past + m0)) {  // tile touches the diagonal / context end
#pragma unroll
        for (int i = 0; i < FA_BN; ++i) {
          const int key = j0 + i;
          if (!(key < T && (!p.causal || key <= pos))) s[i] = -INFINITY;
        }
      }
      float mx0 = s[0], mx1 = s[1], mx2 = s[2], mx3 = s[3];
#pragma unroll
      for (int i = 4; i < FA_BN; i += 4) {
        mx0 = fmaxf(mx0, s[i]); mx1 = fmaxf(mx1, s[i + 1]); mx2 = fmaxf(mx2, s[i + 2]); mx3 = fmaxf(mx3, s[i + 3]);
      }
      const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * p.scale_log2;  // scale > 0: max commutes with it
      float corr = 1.f;
      if (mx > m_ref + FA_RESCALE_THRESHOLD || m_ref == -INFINITY) {  // move the reference maximum
        const float m_new = fmaxf(mx, m_ref);
        if (m_new > -INFINITY) {
          corr = m_ref == -INFINITY ? 1.f : exp2f(m_ref - m_new);
          l *= corr;
          m_ref = m_new;
        }
      }
      const float neg_mref = m_ref == -INFINITY ? 0.f : -m_ref;
      uint32_t pk[FA_BN / 2];
      float l0 = 0.f, l1 = 0.f;
#pragma unroll
      for (int i = 0; i < FA_BN; i += 2) {  // one FFMA + one MUFU.EX2 per score
        float p0, p1;
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p0) : "f"(fmaf(s[i], p.scale_log2, neg_mref)));
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p1) : "f"(fmaf(s[i + 1], p.scale_log2, neg_mref)));
        l0 += p0;
        l1 += p1;
        pk[i / 2] = pack_bf16(p0, p1);
      }
      l += l0 + l1;
      FA_TICK(2);

      if (j > 0) {
        mbar_wait(pv_done, (j - 1) & 1, 20);  // P buffer free, O holds tiles 0..j-1
        tcgen05_fence_after();
        if (__any_sync(0xffffffffu, corr != 1.f)) {  // rare: rescale this warp's 32 rows of O in TMEM
#pragma unroll 1
          for (int c0 = 0; c0 < FA_HD; c0 += 16) {
            tmem_ld16(tO + lane_addr + c0, v);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) * corr);
            tmem_st16(tO + lane_addr + c0, v);
          }
          tmem_st_wait();
        }
      }
      FA_TICK(3);
      // P row -> smem, K-major 128B-swizzled: 16-byte chunk c of row r lives at chunk (c ^ (r & 7))
      uint8_t* prow = fa_smem_raw + (sP - smem_u32(fa_smem_raw)) + row * 128;
#pragma unroll
      for (int c = 0; c < 8; ++c)
        *reinterpret_cast<uint4*>(prow + ((c ^ (row & 7)) << 4)) = make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
      fence_proxy_async_smem();
      tcgen05_fence_before();
      mbar_arrive(p_full);
      FA_TICK(4);
    }
    if (dbg_on) {
      for (int i = 0; i < 5; ++i) p.dbg[i] = tc[i];
      p.dbg[5] = n_tiles;
    }

    // ---- finalize: O / l -> bf16 ----
    mbar_wait(pv_done, (n_tiles - 1) & 1);
    tcgen05_fence_after();
    const float inv = l > 0.f ? 1.f / l : 0.f;
    __nv_bfloat16* dst = p.out + (((long long)b * p.q_len + qi) * p.H + h) * FA_HD;
#pragma unroll 1
    for (int c0 = 0; c0 < FA_HD; c0 += 16) {
      tmem_ld16(tO + lane_addr + c0, v);
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
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(FA_TMEM_COLS) : "memory");
  }
}

bool prefill_attn_tc_supported(const KvView& kv, int hd) {
  static int legacy = -1;
  if (legacy < 0) {
    const char* e = exp_env("DLI_ATTN_LEGACY");
    legacy = (e && e[0] == '1') ? 1 : 0;
  }
  return !legacy && hd == FA_HD && kv.page_tokens >= 8 && kv.page_tokens % 8 == 0 && FA_BN % kv.page_tokens == 0;
}

int launch_prefill_attention_tc(const __nv_bfloat16* q, __nv_bfloat16* out, const KvView& kv, int batch, int q_len, int H,
                                int KV, bool causal, cudaStream_t st) {
  CUtensorMap tq, tk, tv;
  int rc = make_tmap_bf16(&tq, q, (long long)batch * q_len, (long long)H * FA_HD, (long long)H * FA_HD, FA_BM);
  if (rc) return rc;
  const long long pool_rows = (long long)kv.n_pages * KV * kv.page_tokens;
  if ((rc = make_tmap_bf16(&tk, kv.k, pool_rows, FA_HD, FA_HD, kv.page_tokens))) return rc;
  if ((rc = make_tmap_bf16(&tv, kv.v, pool_rows, FA_HD, FA_HD, kv.page_tokens))) return rc;
  DLI_ONCE_PER_DEVICE(cudaFuncSetAttribute(prefill_attn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FA_SMEM));  // > 48 KB dynamic shared memory: opt-in per device
  FaParams p{};
  p.out = out;
  p.block_table = kv.block_table;
  p.seq_lens = kv.seq_lens;
  p.max_pages = kv.max_pages;
  p.page_tokens = kv.page_tokens;
  p.q_len = q_len;
  p.H = H;
  p.KV = KV;
  p.causal = causal ? 1 : 0;
  p.scale_log2 = 1.4426950408889634f / sqrtf((float)FA_HD);
  {
    const char* e = exp_env("DLI_FA_DBG_PTR");  // experiments: device pointer (hex) receiving cycle counters
    p.dbg = e ? reinterpret_cast<unsigned long long*>(strtoull(e, nullptr, 16)) : nullptr;
  }
  dim3 grid((q_len + FA_BM - 1) / FA_BM, H, batch);
  return launch_kernel("prefill_attn_tc_kernel", prefill_attn_tc_kernel, grid, dim3(FA_THREADS), (size_t)FA_SMEM, st, tq, tk, tv, p);
}

}  // namespace dli
