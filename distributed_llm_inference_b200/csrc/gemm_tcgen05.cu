// tcgen05 / TMEM GEMM fed by TMA:   out[m, n] (+)= sum_k x[m, k] * w[n, k]      (bf16 in, fp32 accumulate)
//
// Replaces the seven nn.Linear calls of a reference decoder layer (q/k/v_proj modules.py:61-63, o_proj
// modules.py:112, LlamaMLP gate/up/down modules.py:181) together with the residual adds of
// modules.py:173,182 and the silu(gate)*up product, which live in the epilogue.
//
// One kernel, two operand assignments:
//   * token-major (prefill, m > 128):  A = 128 tokens x K,  B = BN features x K   -> D[token lane, feature col]
//   * swap-AB     (decode,  m <= 128): A = 128 weight rows x K, B = BN (>= batch) tokens x K
//                                                                            -> D[feature lane, token col]
//     so that the weight matrix is the 128-row MMA operand and streams through the tensor core at HBM
//     rate; split-K over gridDim.z fills all SMs and the fp32 partial sums meet in L2 (red.add).
// Both operands are K-major in global memory (nn.Linear stores [out, in]; activations are [tokens, in]),
// so neither needs a transpose: TMA lands 64-element (128-byte) K slices with the 128B swizzle and the
// UMMA shared-memory descriptors consume them directly.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer (both stay converged and
// issue from one elected lane, see elect_one), warps 2-5 = epilogue (warp w reads TMEM lanes 32*(w%4) .. +31, as the
// hardware requires).  Pipelines: smem full/empty mbarrier ring (TMA <-> MMA), tmem_full mbarrier (MMA -> epilogue).
// Three kernels: gemm_tcgen05_kernel (decode: swap-AB, split-K; also the non-persistent token-major form),
// gemm_tcgen05_persistent_kernel (prefill, one CTA per SM, double-buffered accumulators) and
// gemm_tcgen05_pair_kernel (prefill, CTA pairs with cta_group::2 -- the default when m >= 256).
#include "common.cuh"
#include "tc_ptx.cuh"

#include <stdlib.h>

namespace dli {

constexpr int A_TILE_BYTES = BM * BK * 2;
constexpr int GEMM_THREADS = 192;
constexpr int MAX_STAGES = 8;

struct GemmParams {
  void* out;
  long long ldo;
  int m_tokens, n_features;
  int k_blocks, kb_per_split;
  int stages;
  int up_row_offset;  // SILU epilogue: row of w where up_proj starts (= n_features)
  RopeArgs rope;      // QKV_ROPE epilogue only
};

__device__ __forceinline__ float silu(float x) { return x / (1.0f + __expf(-x)); }

// Token-major epilogue of one accumulator tile: lane = token (row arow), columns = features b_row0 ..
template <int BN, int EPI>
__device__ __forceinline__ void epilogue_token_major(uint32_t taddr, int arow, int b_row0, const GemmParams& p) {
  uint32_t v[16];
  if (EPI == DLI_EPI_SILU_MUL_BF16) {
  // lane = token; columns [0, BN/2) = gate, [BN/2, BN) = up of features b_row0 ..
  const bool t_ok = arow < p.m_tokens;
  uint32_t u[16];
#pragma unroll 1
  for (int c0 = 0; c0 < BN / 2; c0 += 16) {
    tmem_ld16(taddr + c0, v);
    tmem_ld16(taddr + BN / 2 + c0, u);
    tmem_ld_wait();
    const int f0 = b_row0 + c0;
    if (t_ok && f0 < p.n_features) {
      __nv_bfloat16* o = (__nv_bfloat16*)p.out + (long long)arow * p.ldo + f0;
      if (f0 + 16 <= p.n_features) {
        uint32_t w[8];
#pragma unroll
        for (int j = 0; j < 8; ++j)
          w[j] = pack_bf16(silu(__uint_as_float(v[2 * j])) * __uint_as_float(u[2 * j]),
                           silu(__uint_as_float(v[2 * j + 1])) * __uint_as_float(u[2 * j + 1]));
        *reinterpret_cast<uint4*>(o) = make_uint4(w[0], w[1], w[2], w[3]);
        *reinterpret_cast<uint4*>(o + 8) = make_uint4(w[4], w[5], w[6], w[7]);
      } else {
        for (int j = 0; j < 16 && f0 + j < p.n_features; ++j)
          o[j] = __float2bfloat16_rn(silu(__uint_as_float(v[j])) * __uint_as_float(u[j]));
      }
    }
  }
} else {
  // lane = token, column = feature
  const bool t_ok = arow < p.m_tokens;
#pragma unroll 1
  for (int c0 = 0; c0 < BN; c0 += 16) {
    tmem_ld16(taddr + c0, v);
    tmem_ld_wait();
    const int f0 = b_row0 + c0;
    if (t_ok && f0 < p.n_features) {
      const bool full = f0 + 16 <= p.n_features;
      if (EPI == DLI_EPI_STORE_BF16) {
        __nv_bfloat16* o = (__nv_bfloat16*)p.out + (long long)arow * p.ldo + f0;
        if (full) {
          uint32_t w[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) w[j] = pack_bf16(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
          *reinterpret_cast<uint4*>(o) = make_uint4(w[0], w[1], w[2], w[3]);
          *reinterpret_cast<uint4*>(o + 8) = make_uint4(w[4], w[5], w[6], w[7]);
        } else {
          for (int j = 0; j < 16 && f0 + j < p.n_features; ++j) o[j] = __float2bfloat16_rn(__uint_as_float(v[j]));
        }
      } else {
        float* o = (float*)p.out + (long long)arow * p.ldo + f0;
        if (full) {
#pragma unroll
          for (int j = 0; j < 16; j += 4) {
            if (EPI == DLI_EPI_ADD_F32) {
              asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(o + j), "f"(__uint_as_float(v[j])),
                           "f"(__uint_as_float(v[j + 1])), "f"(__uint_as_float(v[j + 2])),
                           "f"(__uint_as_float(v[j + 3]))
                           : "memory");
            } else {
              *reinterpret_cast<float4*>(o + j) =
                  make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                              __uint_as_float(v[j + 3]));
            }
          }
        } else {
          for (int j = 0; j < 16 && f0 + j < p.n_features; ++j) {
            if (EPI == DLI_EPI_ADD_F32) atomicAdd(o + j, __uint_as_float(v[j]));
            else o[j] = __uint_as_float(v[j]);
          }
        }
      }
    }
  }
}
}

// Fused epilogue of the q|k|v projection at prefill (token-major tile of 256 features = whole heads): rotates q and
// k with the token's cos/sin rows (modules.py:17-20), writes q densely and k / v into the sequence's page slot
// (cache.py:103-109) — the fp32 projection never goes to HBM and no separate RoPE kernel runs.
template <int BN>
__device__ __forceinline__ void epilogue_qkv_rope(uint32_t taddr, int token, int b_row0, const GemmParams& p) {
  const RopeArgs& r = p.rope;
  const int hd = r.hd, half = hd >> 1;
  const bool t_ok = token < p.m_tokens;
  int l = 0;
  long long slot_off = 0;  // element offset of (page, kv head 0, slot, 0) for this token
  if (t_ok) {
    const int b = token / r.q_len;
    l = token - b * r.q_len;
    const int pos = r.kv.seq_lens[b] + l;
    const int page = r.kv.block_table[(long long)b * r.kv.max_pages + pos / r.kv.page_tokens];
    slot_off = ((long long)page * r.KV * r.kv.page_tokens + pos % r.kv.page_tokens) * hd;
  }
  const long long head_stride = (long long)r.kv.page_tokens * hd;  // between kv heads inside a page
  const float* cs = r.cos + (long long)l * hd;
  const float* sn = r.sin + (long long)l * hd;
  uint32_t lo[16], hi[16];
#pragma unroll 1
  for (int h0 = 0; h0 < BN; h0 += hd) {
    const int f_head = b_row0 + h0;
    if (f_head >= p.n_features) break;  // warp-uniform
    const int head = f_head / hd;
    if (head >= r.H + r.KV) {  // value head: plain conversion
      __nv_bfloat16* dst = r.kv.v + slot_off + (long long)(head - r.H - r.KV) * head_stride;
#pragma unroll 1
      for (int c0 = 0; c0 < hd; c0 += 16) {
        tmem_ld16(taddr + h0 + c0, lo);
        tmem_ld_wait();
        if (t_ok) {
          uint32_t w[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) w[j] = pack_bf16(__uint_as_float(lo[2 * j]), __uint_as_float(lo[2 * j + 1]));
          *reinterpret_cast<uint4*>(dst + c0) = make_uint4(w[0], w[1], w[2], w[3]);
          *reinterpret_cast<uint4*>(dst + c0 + 8) = make_uint4(w[4], w[5], w[6], w[7]);
        }
      }
    } else {
      __nv_bfloat16* dst = head < r.H ? r.q_out + ((long long)token * r.H + head) * hd
                                      : r.kv.k + slot_off + (long long)(head - r.H) * head_stride;
#pragma unroll 1
      for (int c0 = 0; c0 < half; c0 += 16) {
        tmem_ld16(taddr + h0 + c0, lo);
        tmem_ld16(taddr + h0 + half + c0, hi);
        tmem_ld_wait();
        if (t_ok) {
          uint32_t wl[8], wh[8];
#pragma unroll
          for (int j = 0; j < 16; j += 4) {
            const float4 c1 = *reinterpret_cast<const float4*>(cs + c0 + j), c2 = *reinterpret_cast<const float4*>(cs + half + c0 + j);
            const float4 s1 = *reinterpret_cast<const float4*>(sn + c0 + j), s2 = *reinterpret_cast<const float4*>(sn + half + c0 + j);
            const float x1[4] = {__uint_as_float(lo[j]), __uint_as_float(lo[j + 1]), __uint_as_float(lo[j + 2]), __uint_as_float(lo[j + 3])};
            const float x2[4] = {__uint_as_float(hi[j]), __uint_as_float(hi[j + 1]), __uint_as_float(hi[j + 2]), __uint_as_float(hi[j + 3])};
            const float cc1[4] = {c1.x, c1.y, c1.z, c1.w}, cc2[4] = {c2.x, c2.y, c2.z, c2.w};
            const float ss1[4] = {s1.x, s1.y, s1.z, s1.w}, ss2[4] = {s2.x, s2.y, s2.z, s2.w};
            float o1[4], o2[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              o1[e] = x1[e] * cc1[e] - x2[e] * ss1[e];  // rotate_half: first half pairs with -x2
              o2[e] = x2[e] * cc2[e] + x1[e] * ss2[e];
            }
            wl[j / 2] = pack_bf16(o1[0], o1[1]); wl[j / 2 + 1] = pack_bf16(o1[2], o1[3]);
            wh[j / 2] = pack_bf16(o2[0], o2[1]); wh[j / 2 + 1] = pack_bf16(o2[2], o2[3]);
          }
          *reinterpret_cast<uint4*>(dst + c0) = make_uint4(wl[0], wl[1], wl[2], wl[3]);
          *reinterpret_cast<uint4*>(dst + c0 + 8) = make_uint4(wl[4], wl[5], wl[6], wl[7]);
          *reinterpret_cast<uint4*>(dst + half + c0) = make_uint4(wh[0], wh[1], wh[2], wh[3]);
          *reinterpret_cast<uint4*>(dst + half + c0 + 8) = make_uint4(wh[4], wh[5], wh[6], wh[7]);
        }
      }
    }
  }
}

// ---- kernel -------------------------------------------------------------------------------------------
template <int BN, int EPI, bool SWAP>
__global__ void __launch_bounds__(GEMM_THREADS, SWAP ? 2 : 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                    const GemmParams p) {
  constexpr int B_TILE_BYTES = BN * BK * 2;
  constexpr int STAGE_BYTES = A_TILE_BYTES + B_TILE_BYTES;
  constexpr int TMEM_COLS = BN < 32 ? 32 : BN;

  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bars[2 * MAX_STAGES + 1];
  __shared__ uint32_t tmem_base_slot;

  const uint32_t tiles = (smem_u32(smem_raw) + 1023u) & ~1023u;  // 128B swizzle atoms need 1024-byte alignment
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int stages = p.stages;
  const uint32_t full0 = smem_u32(&bars[0]), empty0 = smem_u32(&bars[MAX_STAGES]);
  const uint32_t tmem_full = smem_u32(&bars[2 * MAX_STAGES]);

  const int kb0 = blockIdx.z * p.kb_per_split;
  const int kb1 = min(kb0 + p.kb_per_split, p.k_blocks);
  const int a_row0 = blockIdx.x * ((SWAP && EPI == DLI_EPI_SILU_MUL_BF16) ? BM / 2 : BM);
  const int b_row0 = blockIdx.y * ((!SWAP && EPI == DLI_EPI_SILU_MUL_BF16) ? BN / 2 : BN);

  pdl_trigger();  // the next kernel of the stream may start its own prologue / weight prefetch
  if (threadIdx.x == 0) {
    DLI_TRACE_EV(TR_GEMM + 16 * EPI, 0);
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_b) : "memory");
    for (int s = 0; s < stages; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, 1);
    }
    mbar_init(tmem_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // whole warp: allocate the accumulator columns
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)),
                 "r"((uint32_t)TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_d = tmem_base_slot;

  if (kb1 > kb0) {
    if (warp == 0) {
      {  // ===== TMA producer (converged warp; one elected lane issues: see elect_one) =====
        // The weight operand does not depend on the previous kernel of the stream: under programmatic dependent
        // launch its first `stages` tiles are requested BEFORE griddepcontrol.wait, so weights keep streaming from
        // HBM while the predecessor drains; the activation operand is requested after the wait.
        auto load_weights = [&](int s, int kb) {
          const uint32_t a_dst = tiles + s * STAGE_BYTES, b_dst = a_dst + A_TILE_BYTES;
          if (SWAP && EPI == DLI_EPI_SILU_MUL_BF16) {  // 64 gate rows + 64 up rows of the same 64 features
            tma_load_2d(a_dst, &tma_a, full0 + 8 * s, kb * BK, a_row0);
            tma_load_2d(a_dst + A_TILE_BYTES / 2, &tma_a, full0 + 8 * s, kb * BK, p.up_row_offset + a_row0);
          } else if (SWAP) {
            tma_load_2d(a_dst, &tma_a, full0 + 8 * s, kb * BK, a_row0);
          } else if (EPI == DLI_EPI_SILU_MUL_BF16) {  // gate rows then up rows of the same features
            tma_load_2d(b_dst, &tma_b, full0 + 8 * s, kb * BK, b_row0);
            tma_load_2d(b_dst + B_TILE_BYTES / 2, &tma_b, full0 + 8 * s, kb * BK, p.up_row_offset + b_row0);
          } else {
            tma_load_2d(b_dst, &tma_b, full0 + 8 * s, kb * BK, b_row0);
          }
        };
        auto load_acts = [&](int s, int kb) {
          const uint32_t a_dst = tiles + s * STAGE_BYTES, b_dst = a_dst + A_TILE_BYTES;
          if (SWAP) tma_load_2d(b_dst, &tma_b, full0 + 8 * s, kb * BK, b_row0);
          else tma_load_2d(a_dst, &tma_a, full0 + 8 * s, kb * BK, a_row0);
        };
        const int n_pre = min(stages, kb1 - kb0);
        if (elect_one()) {
          for (int i = 0; i < n_pre; ++i) {  // fresh barriers: every slot is free
            mbar_expect_tx(full0 + 8 * i, STAGE_BYTES);
            load_weights(i, kb0 + i);
          }
        }
        __syncwarp();
        pdl_wait();
        if (lane == 0) DLI_TRACE_EV(TR_GEMM + 16 * EPI, 1);
        if (elect_one()) {
          for (int i = 0; i < n_pre; ++i) load_acts(i, kb0 + i);
        }
        __syncwarp();
        int s = n_pre == stages ? 0 : n_pre;
        uint32_t ph = n_pre == stages ? 1 : 0;
        for (int kb = kb0 + n_pre; kb < kb1; ++kb) {
          mbar_wait(empty0 + 8 * s, ph ^ 1);
          if (elect_one()) {
            mbar_expect_tx(full0 + 8 * s, STAGE_BYTES);
            load_weights(s, kb);
            load_acts(s, kb);
          }
          __syncwarp();
          if (++s == stages) { s = 0; ph ^= 1; }
        }
      }
    } else if (warp == 1) {
      {  // ===== MMA issuer (converged warp; one elected lane issues) =====
        constexpr uint32_t idesc = umma_idesc_bf16(BN);
        int s = 0;
        uint32_t ph = 0;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(full0 + 8 * s, ph);
          tcgen05_fence_after();
          if (elect_one()) {
            const uint32_t a_src = tiles + s * STAGE_BYTES;
            const uint64_t adesc = umma_desc_sw128(a_src), bdesc = umma_desc_sw128(a_src + A_TILE_BYTES);
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) {
              // advancing 16 bf16 (32 bytes) along K inside the swizzle atom: +2 in the (addr >> 4) field
              umma_bf16(tmem_d, adesc + 2 * k, bdesc + 2 * k, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            }
            umma_commit(empty0 + 8 * s);  // frees the smem slot once these MMAs have read it
            if (kb + 1 == kb1) umma_commit(tmem_full);  // accumulator complete
          }
          __syncwarp();
          if (++s == stages) { s = 0; ph ^= 1; }
        }
      }
    } else {  // ===== epilogue: TMEM -> registers -> global =====
      pdl_wait();  // the output buffers belong to the predecessor until it has completed
      mbar_wait(tmem_full, 0, 200);  // 128 threads wait out the whole main loop: poll rarely
      tcgen05_fence_after();
      const int q = warp & 3;  // TMEM lane quarter this warp may read
      const int arow = a_row0 + q * 32 + lane;
      const uint32_t taddr = tmem_d + ((uint32_t)(q * 32) << 16);
      uint32_t v[16];
      if (SWAP) {
        // lane = feature, column = token.  The accumulator is transposed through shared memory (the operand ring
        // is idle once tmem_full has fired) so that global traffic is 128-bit and coalesced along the feature
        // axis: one red.global.add.v4.f32 per 4 features instead of 4 scalar reductions.
        float* stage_f = reinterpret_cast<float*>(smem_raw + (tiles - smem_u32(smem_raw)));  // [BN tokens][128 lanes]
        const int lane_row = q * 32 + lane;
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 16) {
          tmem_ld16(taddr + c0, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) stage_f[(c0 + j) * BM + lane_row] = __uint_as_float(v[j]);
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");  // the four epilogue warps only
        const int e = threadIdx.x - 64;  // 0..127
        if (EPI == DLI_EPI_SILU_MUL_BF16) {
          // lanes 0..63 hold gate, 64..127 up of features a_row0 .. a_row0 + 63; 4 features per thread and token
          for (int idx = e; idx < BN * 16; idx += 128) {
            const int t = idx >> 4, f = (idx & 15) * 4;
            const int tok = b_row0 + t, feat = a_row0 + f;
            if (tok < p.m_tokens && feat < p.n_features) {
              const float4 g4 = *reinterpret_cast<const float4*>(stage_f + t * BM + f);
              const float4 u4 = *reinterpret_cast<const float4*>(stage_f + t * BM + 64 + f);
              __nv_bfloat16* o = (__nv_bfloat16*)p.out + (long long)tok * p.ldo + feat;
              if (feat + 4 <= p.n_features) {
                *reinterpret_cast<uint2*>(o) = make_uint2(pack_bf16(silu(g4.x) * u4.x, silu(g4.y) * u4.y),
                                                          pack_bf16(silu(g4.z) * u4.z, silu(g4.w) * u4.w));
              } else {
                const float gg[4] = {g4.x, g4.y, g4.z, g4.w}, uu[4] = {u4.x, u4.y, u4.z, u4.w};
                for (int j = 0; j < 4 && feat + j < p.n_features; ++j) o[j] = __float2bfloat16_rn(silu(gg[j]) * uu[j]);
              }
            }
          }
        } else {
          for (int idx = e; idx < BN * 32; idx += 128) {
            const int t = idx >> 5, f = (idx & 31) * 4;
            const int tok = b_row0 + t, feat = a_row0 + f;
            if (tok < p.m_tokens && feat < p.n_features) {
              const float4 a4 = *reinterpret_cast<const float4*>(stage_f + t * BM + f);
              const bool full = feat + 4 <= p.n_features;
              if (EPI == DLI_EPI_STORE_BF16) {
                __nv_bfloat16* o = (__nv_bfloat16*)p.out + (long long)tok * p.ldo + feat;
                if (full) {
                  *reinterpret_cast<uint2*>(o) = make_uint2(pack_bf16(a4.x, a4.y), pack_bf16(a4.z, a4.w));
                } else {
                  const float aa[4] = {a4.x, a4.y, a4.z, a4.w};
                  for (int j = 0; j < 4 && feat + j < p.n_features; ++j) o[j] = __float2bfloat16_rn(aa[j]);
                }
              } else {
                float* o = (float*)p.out + (long long)tok * p.ldo + feat;
                if (full && EPI == DLI_EPI_ADD_F32) {
                  asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(o), "f"(a4.x), "f"(a4.y), "f"(a4.z), "f"(a4.w)
                               : "memory");
                } else if (full) {
                  *reinterpret_cast<float4*>(o) = a4;
                } else {
                  const float aa[4] = {a4.x, a4.y, a4.z, a4.w};
                  for (int j = 0; j < 4 && feat + j < p.n_features; ++j) {
                    if (EPI == DLI_EPI_ADD_F32) atomicAdd(o + j, aa[j]);
                    else o[j] = aa[j];
                  }
                }
              }
            }
          }
        }
      } else {
        epilogue_token_major<BN, EPI>(taddr, arow, b_row0, p);
      }
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) DLI_TRACE_EV(TR_GEMM + 16 * EPI, 2);
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"((uint32_t)TMEM_COLS) : "memory");
  }
}

// ---- persistent token-major kernel (prefill) -------------------------------------------------------------
// One CTA per SM loops over output tiles (static round-robin, token tiles fastest so that concurrently running
// CTAs share weight tiles in L2).  The accumulator is double-buffered in TMEM (2 x BN columns), so the epilogue of
// tile i (tcgen05.ld -> global) overlaps the TMA/MMA main loop of tile i+1; the smem operand ring runs straight
// through tile boundaries.  Barriers: smem full/empty per stage, tmem_full/tmem_empty per accumulator.
// Tile rasterisation of the persistent kernel: bands of `group_m` token tiles; inside a band feature tiles advance
// slowest-but-one, so the CTAs running at any moment cover (group_m token tiles) x (a few feature tiles): the band's
// activations stay in L2 for the whole sweep and every weight tile is fetched from HBM once per band (ncu on the
// naive order showed 8x the algorithmic DRAM traffic: the full activation matrix did not survive in L2).
__device__ __forceinline__ void tile_coords(int t, int tiles_m, int tiles_n, int group_m, int& mt, int& nt) {
  const int band_sz = group_m * tiles_n;
  const int band = t / band_sz, r = t - band * band_sz;
  const int gsize = min(group_m, tiles_m - band * group_m);
  nt = r / gsize;
  mt = band * group_m + (r - nt * gsize);
}

template <int BN, int EPI>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_persistent_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                               const GemmParams p, const int tiles_m, const int tiles_n, const int group_m) {
  const int n_tiles = tiles_m * tiles_n;
  constexpr int B_TILE_BYTES = BN * BK * 2;
  constexpr int STAGE_BYTES = A_TILE_BYTES + B_TILE_BYTES;
  constexpr int TMEM_COLS = 2 * BN;  // 256 or 512
  constexpr int N_PER_TILE = EPI == DLI_EPI_SILU_MUL_BF16 ? BN / 2 : BN;  // output features per tile

  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bars[2 * MAX_STAGES + 4];
  __shared__ uint32_t tmem_base_slot;

  const uint32_t tiles = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int stages = p.stages;
  const uint32_t full0 = smem_u32(&bars[0]), empty0 = smem_u32(&bars[MAX_STAGES]);
  const uint32_t tfull0 = smem_u32(&bars[2 * MAX_STAGES]), tempty0 = smem_u32(&bars[2 * MAX_STAGES + 2]);
  const int n_kb = p.k_blocks;

  pdl_trigger();
  if (threadIdx.x == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_b) : "memory");
    for (int s = 0; s < stages; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull0 + 8 * a, 1);
      mbar_init(tempty0 + 8 * a, 4);  // one arrival per epilogue warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)),
                 "r"((uint32_t)TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_d = tmem_base_slot;

  if (warp == 0) {
    {  // ===== TMA producer (converged warp; one elected lane issues) =====
      int s = 0;
      uint32_t ph = 0;
      bool waited = false;
      for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        int mt, nt;
        tile_coords(t, tiles_m, tiles_n, group_m, mt, nt);
        const int a_row0 = mt * BM, b_row0 = nt * N_PER_TILE;
        for (int kb = 0; kb < n_kb; ++kb) {
          mbar_wait(empty0 + 8 * s, ph ^ 1);
          const uint32_t a_dst = tiles + s * STAGE_BYTES, b_dst = a_dst + A_TILE_BYTES;
          if (elect_one()) {
            mbar_expect_tx(full0 + 8 * s, STAGE_BYTES);
            if (EPI == DLI_EPI_SILU_MUL_BF16) {  // gate rows then up rows of the same features
              tma_load_2d(b_dst, &tma_b, full0 + 8 * s, kb * BK, b_row0);
              tma_load_2d(b_dst + B_TILE_BYTES / 2, &tma_b, full0 + 8 * s, kb * BK, p.up_row_offset + b_row0);
            } else {
              tma_load_2d(b_dst, &tma_b, full0 + 8 * s, kb * BK, b_row0);
            }
          }
          __syncwarp();
          if (!waited) {  // weights do not depend on the previous kernel; activations do
            pdl_wait();
            waited = true;
          }
          if (elect_one()) tma_load_2d(a_dst, &tma_a, full0 + 8 * s, kb * BK, a_row0);
          __syncwarp();
          if (++s == stages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    {  // ===== MMA issuer (converged warp; one elected lane issues) =====
      constexpr uint32_t idesc = umma_idesc_bf16(BN);
      int s = 0, it = 0;
      uint32_t ph = 0;
      for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
        const int acc = it & 1;
        mbar_wait(tempty0 + 8 * acc, (((uint32_t)it >> 1) & 1) ^ 1);  // epilogue has drained this accumulator
        tcgen05_fence_after();
        const uint32_t d = tmem_d + (uint32_t)(acc * BN);
        for (int kb = 0; kb < n_kb; ++kb) {
          mbar_wait(full0 + 8 * s, ph);
          tcgen05_fence_after();
          if (elect_one()) {
            const uint32_t a_src = tiles + s * STAGE_BYTES;
            const uint64_t adesc = umma_desc_sw128(a_src), bdesc = umma_desc_sw128(a_src + A_TILE_BYTES);
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) umma_bf16(d, adesc + 2 * k, bdesc + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
            umma_commit(empty0 + 8 * s);
            if (kb + 1 == n_kb) umma_commit(tfull0 + 8 * acc);
          }
          __syncwarp();
          if (++s == stages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else {  // ===== epilogue warps =====
    pdl_wait();
    const int q = warp & 3;
    int it = 0;
    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
      const int acc = it & 1;
      int mt, nt;
      tile_coords(t, tiles_m, tiles_n, group_m, mt, nt);
      const int a_row0 = mt * BM, b_row0 = nt * N_PER_TILE;
      mbar_wait(tfull0 + 8 * acc, ((uint32_t)it >> 1) & 1, 100);
      tcgen05_fence_after();
      const uint32_t taddr = tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN);
      if constexpr (EPI == DLI_EPI_QKV_ROPE_INTERNAL) epilogue_qkv_rope<BN>(taddr, a_row0 + q * 32 + lane, b_row0, p);
      else epilogue_token_major<BN, EPI>(taddr, a_row0 + q * 32 + lane, b_row0, p);
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tempty0 + 8 * acc) : "memory");
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"((uint32_t)TMEM_COLS) : "memory");
  }
}


// ---- CTA-pair persistent kernel (prefill, cta_group::2) ----------------------------------------------------
// The same persistent, double-buffered-accumulator structure as above, but two CTAs on the SMs of one TPC work on
// ONE 256-token x 256-feature tile: each CTA stages its own 128 token rows of A and only HALF of the B tile (128
// feature rows); `tcgen05.mma.cta_group::2` (M = 256), issued by the leader CTA alone, feeds both tensor cores from
// both halves.  Per k-block a CTA now moves 32 KB through TMA and shared memory instead of 48 KB, which is what the
// single-CTA kernel was short of (ncu: tensor pipe 81-83 % active with the operand ring never empty).
//   barriers  full[s]   leader only: 1 arrival (expect_tx of both CTAs' bytes) + the tx of both CTAs' TMA loads
//             empty[s]  in both CTAs: arrival by the leader's multicast tcgen05.commit
//             tfull[a]  in both CTAs: arrival by the leader's multicast commit after a tile's last k-block
//             tempty[a] leader only: 8 arrivals = the 4 epilogue warps of each CTA (the peer's arrive remotely)
// For the fused silu(gate)*up epilogue the pair's B halves are the gate rows (rank 0) and the up rows (rank 1) of
// the same 128 features, so the accumulator columns are [gate | up] exactly as in the single-CTA kernel.
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa_rank(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
// TMA load into this CTA's shared memory whose completion bytes are counted on a barrier of the pair's leader
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* map, uint32_t bar_cluster, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar_cluster), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {  // arrives on `bar` of BOTH CTAs of the pair
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
               "h"((uint16_t)3)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar_cluster) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(bar_cluster) : "memory");
}

constexpr int PAIR_BN = 256;       // accumulator columns per CTA (features of the pair's tile)
constexpr int PAIR_B_ROWS = 128;   // B rows each CTA stages
constexpr int PAIR_STAGE_BYTES = A_TILE_BYTES + PAIR_B_ROWS * BK * 2;  // 32 KB

template <int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_pair_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                         const GemmParams p, const int tiles_m, const int tiles_n, const int group_m) {
  constexpr int BN = PAIR_BN;
  constexpr int TMEM_COLS = 2 * BN;
  constexpr int N_PER_TILE = EPI == DLI_EPI_SILU_MUL_BF16 ? BN / 2 : BN;
  const int n_tiles = tiles_m * tiles_n;

  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bars[2 * MAX_STAGES + 4];
  __shared__ uint32_t tmem_base_slot;

  const uint32_t tiles = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int stages = p.stages;
  const uint32_t full0 = smem_u32(&bars[0]), empty0 = smem_u32(&bars[MAX_STAGES]);
  const uint32_t tfull0 = smem_u32(&bars[2 * MAX_STAGES]), tempty0 = smem_u32(&bars[2 * MAX_STAGES + 2]);
  const int n_kb = p.k_blocks;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  const uint32_t full0_leader = mapa_rank(full0, 0), tempty0_leader = mapa_rank(tempty0, 0);

  pdl_trigger();
  if (threadIdx.x == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_b) : "memory");
    for (int s = 0; s < stages; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull0 + 8 * a, 1);
      mbar_init(tempty0 + 8 * a, 8);  // 4 epilogue warps of each CTA
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // the same warp of both CTAs allocates the pair's accumulator columns
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)),
                 "r"((uint32_t)TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  cluster_sync_all();  // barriers of both CTAs exist before anyone signals them
  tcgen05_fence_after();
  const uint32_t tmem_d = tmem_base_slot;

  if (warp == 0) {
    // ===== TMA producer of this CTA: its 128 token rows of A and its half of B =====
    int s = 0;
    uint32_t ph = 0;
    bool waited = false;
    for (int t = pair; t < n_tiles; t += n_pairs) {
      int mt, nt;
      tile_coords(t, tiles_m, tiles_n, group_m, mt, nt);
      const int a_row0 = mt * 2 * BM + (int)rank * BM;
      const int b_row0 = EPI == DLI_EPI_SILU_MUL_BF16 ? nt * N_PER_TILE + (int)rank * p.up_row_offset
                                                      : nt * N_PER_TILE + (int)rank * PAIR_B_ROWS;
      for (int kb = 0; kb < n_kb; ++kb) {
        mbar_wait(empty0 + 8 * s, ph ^ 1);
        const uint32_t a_dst = tiles + s * PAIR_STAGE_BYTES, b_dst = a_dst + A_TILE_BYTES;
        if (elect_one()) {
          if (leader) mbar_expect_tx(full0 + 8 * s, 2 * PAIR_STAGE_BYTES);
          tma_load_2d_pair(b_dst, &tma_b, full0_leader + 8 * s, kb * BK, b_row0);
        }
        __syncwarp();
        if (!waited) {  // weights do not depend on the previous kernel; activations do
          pdl_wait();
          waited = true;
        }
        if (elect_one()) tma_load_2d_pair(a_dst, &tma_a, full0_leader + 8 * s, kb * BK, a_row0);
        __syncwarp();
        if (++s == stages) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (leader) {
      // ===== MMA issuer of the pair =====
      constexpr uint32_t idesc =
          (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((2 * BM) >> 4) << 24);
      int s = 0, it = 0;
      uint32_t ph = 0;
      for (int t = pair; t < n_tiles; t += n_pairs, ++it) {
        const int acc = it & 1;
        mbar_wait(tempty0 + 8 * acc, (((uint32_t)it >> 1) & 1) ^ 1);  // both CTAs have drained this accumulator
        tcgen05_fence_after();
        const uint32_t d = tmem_d + (uint32_t)(acc * BN);
        for (int kb = 0; kb < n_kb; ++kb) {
          mbar_wait(full0 + 8 * s, ph);
          tcgen05_fence_after();
          if (elect_one()) {
            const uint32_t a_src = tiles + s * PAIR_STAGE_BYTES;
            const uint64_t adesc = umma_desc_sw128(a_src), bdesc = umma_desc_sw128(a_src + A_TILE_BYTES);
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k)
              umma_bf16_pair(d, adesc + 2 * k, bdesc + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
            umma_commit_pair(empty0 + 8 * s);
            if (kb + 1 == n_kb) umma_commit_pair(tfull0 + 8 * acc);
          }
          __syncwarp();
          if (++s == stages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else {  // ===== epilogue warps: this CTA's 128 token rows of the tile =====
    pdl_wait();
    const int q = warp & 3;
    int it = 0;
    for (int t = pair; t < n_tiles; t += n_pairs, ++it) {
      const int acc = it & 1;
      int mt, nt;
      tile_coords(t, tiles_m, tiles_n, group_m, mt, nt);
      const int a_row0 = mt * 2 * BM + (int)rank * BM, b_row0 = nt * N_PER_TILE;
      mbar_wait(tfull0 + 8 * acc, ((uint32_t)it >> 1) & 1, 100);
      tcgen05_fence_after();
      const uint32_t taddr = tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN);
      if constexpr (EPI == DLI_EPI_QKV_ROPE_INTERNAL) epilogue_qkv_rope<BN>(taddr, a_row0 + q * 32 + lane, b_row0, p);
      else epilogue_token_major<BN, EPI>(taddr, a_row0 + q * 32 + lane, b_row0, p);
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(tempty0_leader + 8 * acc);
    }
  }

  tcgen05_fence_before();
  cluster_sync_all();  // neither CTA leaves (or frees TMEM) while its peer may still read its smem / signal its barriers
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"((uint32_t)TMEM_COLS) : "memory");
  }
}


// ---- host side ----------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

// 2-D bf16 tensor [rows, cols] with row pitch ld elements; box = [box_rows, 64 cols], 128B swizzle.
int make_tmap_bf16(CUtensorMap* map, const void* base, long long rows, long long cols, long long ld, int box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return fail(DLI_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (ld % 8) != 0)
    return fail(DLI_ERR_ARG, "TMA operand needs a 16-byte aligned base and a row pitch that is a multiple of 8 elements");
  if (box_rows < 1 || box_rows > 256) return fail(DLI_ERR_ARG, "TMA box rows %d out of range", box_rows);
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(DLI_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return DLI_OK;
}

template <int BN, int EPI, bool SWAP>
static int launch_variant(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, dim3 grid, cudaStream_t st) {
  constexpr int STAGE_BYTES = A_TILE_BYTES + BN * BK * 2;
  const size_t smem = (size_t)p.stages * STAGE_BYTES + 1024;
  auto kern = gemm_tcgen05_kernel<BN, EPI, SWAP>;
  DLI_ONCE_PER_DEVICE(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 1024));  // > 48 KB dynamic shared memory: opt-in per device
  return launch_kernel("gemm_tcgen05_kernel", kern, grid, dim3(GEMM_THREADS), smem, st, ta, tb, p);
}

template <int BN, int EPI>
static int launch_persistent(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int tiles_m, int tiles_n,
                             cudaStream_t st) {
  constexpr int STAGE_BYTES = A_TILE_BYTES + BN * BK * 2;
  const size_t smem = (size_t)p.stages * STAGE_BYTES + 1024;
  auto kern = gemm_tcgen05_persistent_kernel<BN, EPI>;
  DLI_ONCE_PER_DEVICE(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 1024));  // > 48 KB dynamic shared memory: opt-in per device
  const int n_tiles = tiles_m * tiles_n;
  const int grid = n_tiles < sm_count() ? n_tiles : sm_count();
  static int group_env = -1;
  if (group_env < 0) {
    const char* e = exp_env("DLI_GROUP_M");
    group_env = e ? atoi(e) : 0;
  }
  // band height: activations of a band (group_m x 128 tokens x K) should fit comfortably in one die's share of L2
  int group_m = group_env > 0 ? group_env : 16;
  if (group_m > tiles_m) group_m = tiles_m;
  return launch_kernel("gemm_tcgen05_persistent_kernel", kern, dim3(grid), dim3(GEMM_THREADS), smem, st, ta, tb, p, tiles_m,
                       tiles_n, group_m);
}


static bool pair_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = exp_env("DLI_NO_PAIR");
    v = (e && e[0] == '1') ? 0 : 1;
  }
  return v == 1;
}

// CTA-pair kernel: tiles_m counts 256-token tiles; the grid is the largest even number of CTAs <= #SMs
template <int EPI>
static int launch_pair(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p_in, int tiles_m, int tiles_n,
                       cudaStream_t st) {
  GemmParams p = p_in;
  p.stages = std::min(6, std::max(2, p.k_blocks));
  const size_t smem = (size_t)p.stages * PAIR_STAGE_BYTES + 1024;
  auto kern = gemm_tcgen05_pair_kernel<EPI>;
  DLI_ONCE_PER_DEVICE(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 1024));  // > 48 KB dynamic shared memory: opt-in per device
  const int n_tiles = tiles_m * tiles_n;
  const int pairs = std::min(n_tiles, sm_count() / 2);
  int group_m = 8;  // bands of 8 x 256 tokens: the same 2048-token activation band as the single-CTA kernel
  {
    static int group_env = -1;
    if (group_env < 0) {
      const char* e = exp_env("DLI_GROUP_M");
      group_env = e ? atoi(e) : 0;
    }
    if (group_env > 0) group_m = std::max(1, group_env / 2);
  }
  if (group_m > tiles_m) group_m = tiles_m;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(2 * pairs);
  cfg.blockDim = dim3(GEMM_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = (pdl_enabled() && !pdl_take_barrier()) ? 1 : 0;  // cluster dims are compiled into the kernel
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, ta, tb, p, tiles_m, tiles_n, group_m);
  if (e != cudaSuccess) return fail(DLI_ERR_CUDA, "launch of gemm_tcgen05_pair_kernel failed: %s", cudaGetErrorString(e));
  count_launch();
  return DLI_OK;
}

template <int BN>
static int launch_persistent_epi(int epi, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int tiles_m,
                                 int tiles_n, cudaStream_t st) {
  switch (epi) {
    case DLI_EPI_STORE_F32: return launch_persistent<BN, DLI_EPI_STORE_F32>(ta, tb, p, tiles_m, tiles_n, st);
    case DLI_EPI_ADD_F32: return launch_persistent<BN, DLI_EPI_ADD_F32>(ta, tb, p, tiles_m, tiles_n, st);
    case DLI_EPI_STORE_BF16: return launch_persistent<BN, DLI_EPI_STORE_BF16>(ta, tb, p, tiles_m, tiles_n, st);
    default: return fail(DLI_ERR_ARG, "unknown epilogue %d", epi);
  }
}

static bool persistent_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = exp_env("DLI_NO_PERSISTENT");
    v = (e && e[0] == '1') ? 0 : 1;
  }
  return v == 1;
}

bool gemm_persistent_enabled() { return persistent_enabled(); }

template <int BN, bool SWAP>
static int launch_epi(int epi, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, dim3 grid, cudaStream_t st) {
  switch (epi) {
    case DLI_EPI_STORE_F32: return launch_variant<BN, DLI_EPI_STORE_F32, SWAP>(ta, tb, p, grid, st);
    case DLI_EPI_ADD_F32: return launch_variant<BN, DLI_EPI_ADD_F32, SWAP>(ta, tb, p, grid, st);
    case DLI_EPI_STORE_BF16: return launch_variant<BN, DLI_EPI_STORE_BF16, SWAP>(ta, tb, p, grid, st);
    default: return fail(DLI_ERR_ARG, "unknown epilogue %d", epi);
  }
}

GemmPlan plan_gemm(int m, int n, int k, int epilogue) {
  GemmPlan g{};
  g.k_blocks = (k + BK - 1) / BK;
  const int sms = sm_count();
  if (m <= 128 && epilogue != DLI_EPI_QKV_ROPE_INTERNAL) {
    const bool fused_silu = epilogue == DLI_EPI_SILU_MUL_BF16;  // 64 gate + 64 up rows per CTA, never split
    g.swap = true;
    g.bn = m <= 16 ? 16 : m <= 32 ? 32 : m <= 64 ? 64 : 128;
    g.a_box_rows = fused_silu ? BM / 2 : BM;  // weights
    g.b_box_rows = g.bn;                      // tokens
    const int tiles = fused_silu ? (n + BM / 2 - 1) / (BM / 2) : (n + BM - 1) / BM;
    int splits = 1;
    if (epilogue == DLI_EPI_ADD_F32) {
      // two co-resident CTAs per SM keep ~200 KB of weight tiles in flight per SM; every split keeps >= 4 k-blocks
      static int min_kb = 0, ctas_per_sm_x2 = 0;
      if (!min_kb) {
        const char* e = exp_env("DLI_MIN_KB");
        min_kb = e ? max(1, atoi(e)) : 4;
        const char* f = exp_env("DLI_FILL_X2");
        ctas_per_sm_x2 = f ? max(1, atoi(f)) : 4;  // target CTAs per SM, times two
      }
      const int want = (ctas_per_sm_x2 * sms / 2) / tiles;
      splits = max(1, min(want, g.k_blocks / min_kb));
      splits = min(splits, 32);
    }
    g.kb_per_split = (g.k_blocks + splits - 1) / splits;
    g.splits = (g.k_blocks + g.kb_per_split - 1) / g.kb_per_split;
    const int stage_bytes = A_TILE_BYTES + g.bn * BK * 2;
    g.stages = min(MAX_STAGES, (112 * 1024 - 2048) / stage_bytes);
    g.stages = max(2, min(g.stages, g.kb_per_split));
    g.grid_x = tiles;
    g.grid_y = (m + g.bn - 1) / g.bn;
  } else {
    g.swap = false;
    const int tm = (m + BM - 1) / BM;
    if (epilogue == DLI_EPI_SILU_MUL_BF16) {
      g.bn = 256;               // 128 gate + 128 up columns
      g.b_box_rows = 128;
      g.grid_y = (n + 127) / 128;
    } else {
      auto waves = [&](int bn) { return (double)(((long long)tm * ((n + bn - 1) / bn) + sms - 1) / sms) * bn; };
      g.bn = (n % 256 == 0 && waves(256) * 0.95 <= waves(128)) ? 256 : 128;
      if (epilogue == DLI_EPI_QKV_ROPE_INTERNAL) g.bn = 256;  // whole heads per tile
      g.b_box_rows = g.bn;
      g.grid_y = (n + g.bn - 1) / g.bn;
    }
    g.a_box_rows = BM;
    // CTA pairs (cta_group::2): each CTA stages 128 weight rows of a 256-column tile (gate rows / up rows for SILU)
    g.pair = persistent_enabled() && pair_enabled() && m >= 2 * BM &&
             (epilogue == DLI_EPI_SILU_MUL_BF16 ? n % 128 == 0 : (g.bn == 256 && n % 256 == 0));
    if (g.pair) g.b_box_rows = 128;
    g.grid_x = tm;
    g.splits = 1;
    g.kb_per_split = g.k_blocks;
    const int stage_bytes = A_TILE_BYTES + g.bn * BK * 2;
    g.stages = min(MAX_STAGES, (226 * 1024 - 2048) / stage_bytes);
    g.stages = max(2, min(g.stages, g.k_blocks));
  }
  return g;
}

int launch_gemm_planned(const GemmPlan& g, const CUtensorMap& tmap_x, const CUtensorMap& tmap_w, void* out, long long ldo,
                        int m, int n, int epilogue, cudaStream_t st, const RopeArgs* rope) {
  GemmParams p{};
  if (epilogue == DLI_EPI_QKV_ROPE_INTERNAL) {
    if (!rope || g.swap || g.bn != 256 || n % 256 != 0 || 256 % rope->hd != 0 || !persistent_enabled())
      return fail(DLI_ERR_ARG, "fused qkv+rope epilogue needs the persistent 256-column path");
    p.rope = *rope;
  }
  p.out = out;
  p.ldo = ldo;
  p.m_tokens = m;
  p.n_features = n;
  p.k_blocks = g.k_blocks;
  p.kb_per_split = g.kb_per_split;
  p.stages = g.stages;
  p.up_row_offset = n;
  dim3 grid(g.grid_x, g.grid_y, g.splits);
  if (g.swap && epilogue == DLI_EPI_SILU_MUL_BF16) {
    switch (g.bn) {
      case 16: return launch_variant<16, DLI_EPI_SILU_MUL_BF16, true>(tmap_w, tmap_x, p, grid, st);
      case 32: return launch_variant<32, DLI_EPI_SILU_MUL_BF16, true>(tmap_w, tmap_x, p, grid, st);
      case 64: return launch_variant<64, DLI_EPI_SILU_MUL_BF16, true>(tmap_w, tmap_x, p, grid, st);
      default: return launch_variant<128, DLI_EPI_SILU_MUL_BF16, true>(tmap_w, tmap_x, p, grid, st);
    }
  }
  if (g.swap) {
    // A = weights, B = tokens
    switch (g.bn) {
      case 16: return launch_epi<16, true>(epilogue, tmap_w, tmap_x, p, grid, st);
      case 32: return launch_epi<32, true>(epilogue, tmap_w, tmap_x, p, grid, st);
      case 64: return launch_epi<64, true>(epilogue, tmap_w, tmap_x, p, grid, st);
      default: return launch_epi<128, true>(epilogue, tmap_w, tmap_x, p, grid, st);
    }
  }
  if (persistent_enabled() && g.pair) {  // prefill, CTA pairs: 256-token x 256-feature tiles, B halves shared
    const int tm2 = (m + 2 * BM - 1) / (2 * BM);
    switch (epilogue) {
      case DLI_EPI_QKV_ROPE_INTERNAL: return launch_pair<DLI_EPI_QKV_ROPE_INTERNAL>(tmap_x, tmap_w, p, tm2, g.grid_y, st);
      case DLI_EPI_SILU_MUL_BF16: return launch_pair<DLI_EPI_SILU_MUL_BF16>(tmap_x, tmap_w, p, tm2, g.grid_y, st);
      case DLI_EPI_STORE_F32: return launch_pair<DLI_EPI_STORE_F32>(tmap_x, tmap_w, p, tm2, g.grid_y, st);
      case DLI_EPI_ADD_F32: return launch_pair<DLI_EPI_ADD_F32>(tmap_x, tmap_w, p, tm2, g.grid_y, st);
      case DLI_EPI_STORE_BF16: return launch_pair<DLI_EPI_STORE_BF16>(tmap_x, tmap_w, p, tm2, g.grid_y, st);
      default: return fail(DLI_ERR_ARG, "unknown epilogue %d", epilogue);
    }
  }
  if (persistent_enabled()) {  // prefill: persistent CTAs, double-buffered TMEM accumulators
    if (epilogue == DLI_EPI_QKV_ROPE_INTERNAL)
      return launch_persistent<256, DLI_EPI_QKV_ROPE_INTERNAL>(tmap_x, tmap_w, p, g.grid_x, g.grid_y, st);
    if (epilogue == DLI_EPI_SILU_MUL_BF16)
      return launch_persistent<256, DLI_EPI_SILU_MUL_BF16>(tmap_x, tmap_w, p, g.grid_x, g.grid_y, st);
    if (g.bn == 256) return launch_persistent_epi<256>(epilogue, tmap_x, tmap_w, p, g.grid_x, g.grid_y, st);
    return launch_persistent_epi<128>(epilogue, tmap_x, tmap_w, p, g.grid_x, g.grid_y, st);
  }
  if (epilogue == DLI_EPI_SILU_MUL_BF16) return launch_variant<256, DLI_EPI_SILU_MUL_BF16, false>(tmap_x, tmap_w, p, grid, st);
  if (g.bn == 256) return launch_epi<256, false>(epilogue, tmap_x, tmap_w, p, grid, st);
  return launch_epi<128, false>(epilogue, tmap_x, tmap_w, p, grid, st);
}

int launch_gemm(const void* x, long long ldx, const void* w, void* out, long long ldo, int m, int n, int k, int epilogue,
                cudaStream_t st) {
  DLI_REQUIRE(x && w && out, "dli_gemm: null pointer");
  DLI_REQUIRE(m > 0 && n > 0 && k > 0, "dli_gemm: empty problem (m=%d n=%d k=%d)", m, n, k);
  DLI_REQUIRE(k % 8 == 0 && ldx % 8 == 0, "dli_gemm: k and ldx must be multiples of 8 (k=%d ldx=%lld)", k, ldx);
  DLI_REQUIRE(epilogue >= 0 && epilogue <= DLI_EPI_SILU_MUL_BF16, "dli_gemm: unknown epilogue %d", epilogue);
  const bool f32_out = epilogue == DLI_EPI_STORE_F32 || epilogue == DLI_EPI_ADD_F32;
  DLI_REQUIRE((reinterpret_cast<uintptr_t>(out) & 15) == 0 && ldo % (f32_out ? 4 : 8) == 0,
              "dli_gemm: output must be 16-byte aligned with ldo a multiple of %d", f32_out ? 4 : 8);
  int epi = epilogue;
  GemmPlan g = plan_gemm(m, n, k, epi);
  if (g.swap && epi == DLI_EPI_STORE_F32) {
    // weight-streaming path accumulates split-K partials in L2: clear the destination, then add
    DLI_CUDA_OK(cudaMemset2DAsync(out, (size_t)ldo * 4, 0, (size_t)n * 4, (size_t)m, st));
    count_launch();
    epi = DLI_EPI_ADD_F32;
    g = plan_gemm(m, n, k, epi);
  }
  const long long w_rows = epilogue == DLI_EPI_SILU_MUL_BF16 ? 2LL * n : n;
  CUtensorMap tx, tw;
  int rc = make_tmap_bf16(&tx, x, m, k, ldx, g.swap ? g.b_box_rows : g.a_box_rows);
  if (rc) return rc;
  rc = make_tmap_bf16(&tw, w, w_rows, k, k, g.swap ? g.a_box_rows : g.b_box_rows);
  if (rc) return rc;
  return launch_gemm_planned(g, tx, tw, out, ldo, m, n, epi, st);
}

DLI_TRACE_SETTER(trace_set_gemm)

}  // namespace dli
