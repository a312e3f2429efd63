// Weight-only int8 decode GEMM:  out[m, n] (+)= scale[n] * sum_k x[m, k] * (q[n, k] - 128)      (m <= 128)
//
// The reference's only "optimised" mode turns every nn.Linear into bitsandbytes Linear8bitLt with
// has_fp16_weights=False (utils/model.py:93-123): weights are stored as row-wise absmax int8 (one fp32 scale per output
// row).  Decode is bound by the bytes of the weights, so storing them in 8 bits halves the step's HBM traffic.  This
// kernel keeps that weight format (biased uint8 = int8 + 128, scale = absmax / 127) and the 16-bit activations of the
// bf16 path: the int8 tile that TMA lands in shared memory is widened to bf16 -- exactly, every integer in
// [-128, 127] is a bf16 -- by eight more warps of the CTA and handed to the tensor core THROUGH TENSOR MEMORY: each
// converter thread owns one weight row (= one TMEM lane), reads its 32 int8 from shared memory, and tcgen05.st's the
// 16 bf16 pairs into the A-operand slot; the MMA is the TS form of the bf16 one (A from TMEM, activations from shared
// memory, fp32 accumulators in TMEM) and the row scale is applied to the accumulator in the epilogue.  There is no
// activation quantisation and therefore no outlier decomposition.
//
// Same shape as the bf16 weight-streaming kernel (gemm_tcgen05.cu, swap-AB: 128 weight rows are the MMA-M operand,
// the batch is MMA-N; split-K over gridDim.z; TMA producer warp, MMA warp), with one more hop per k-block:
//   raw_full[r]   TMA (int8 weights + bf16 activations of k-block kb into raw stage r = kb mod R) -> converter warps
//   a_full[c]     converter warps (bf16 weights in TMEM slot c = kb mod 4, tcgen05.wait::st) -> MMA warp
//   a_empty[c]    tcgen05.commit -> converter warps;   raw_empty[r]  tcgen05.commit -> TMA producer (the activation
//                 tile the MMA reads lives in the raw stage)
// First cut (kept in git history): widened tiles written back to shared memory for an SS-mode MMA.  That needs a
// fence.proxy.async per k-block, which in a main loop with a deep ring of TMA loads in flight cost ~1000 cycles per
// k-block (per-phase counters, tests/exp_w8_timing.py) -- the kernel streamed int8 no faster than the bf16 kernel
// streams bf16.  Tensor memory needs no proxy fence.  The raw tile uses the 128-byte TMA swizzle so that 32 lanes reading
// 32 different rows hit all shared-memory banks.  A raw stage covers two k-blocks (128 int8 per row): with 64-byte rows the
// kernel was bound by the NUMBER of row requests TMA issues, not by bytes (converters waited ~1470 cycles per k-block for
// data while the widening took 200), i.e. it streamed int8 rows no faster than the bf16 kernel streams bf16 rows.
#include "common.cuh"
#include "tc_ptx.cuh"

#include <algorithm>
#include <stdlib.h>

namespace dli {

// NCW = 8 or 16 converter warps widen the int8 tiles (and write the output afterwards), + TMA producer warp + MMA warp.
// A CTA walks its k-blocks one after the other and the widening of a k-block is a dependent chain per thread, so a grid
// with few, long CTAs (gate|up: one CTA per 64 features, never split, all of K) is bound by that chain: it gets 16 warps,
// each thread a quarter of a row per k-block instead of half.
constexpr int W8_RAW_K = 2 * BK;           // a raw stage holds TWO k-blocks: 128-byte rows, like the bf16 kernel's TMA rows
constexpr int W8_RAW_BYTES = BM * W8_RAW_K;  // 128 rows x 128 int8
#ifndef W8_TMAP_ELT
#define W8_TMAP_ELT 1   // bytes per element the weight tensor map is described in (experiments: 4 = uint32 view)
#endif
constexpr int W8_MAX_STAGES = 5;   // raw ring: int8 weight tile + two activation tiles per stage (what is in flight from HBM)
constexpr int W8_A_SLOTS = 4;      // widened bf16 weight tiles (32 TMEM columns each) between converters and MMA warp
constexpr int W8_A_COLS = BK / 2;  // 64 bf16 = 32 columns of bf16 pairs

// D[tmem] (+)= A[tmem] * B[smem]: A = 128 lanes x 16 bf16 (8 columns of bf16 pairs), B K-major in shared memory
__device__ __forceinline__ void w8_umma_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

struct W8Params {
  void* out;
  long long ldo;
  int m_tokens, n_features;
  int k_blocks, kb_per_split;
  int stages;
  int up_row_offset;   // SILU epilogue: row of q / scale where up_proj starts (= n_features)
  const float* scale;  // [rows of q]
  long long* dbg;      // W8_TIMING builds: per-phase cycle counters of CTA 0 (experiments)
};

__device__ __forceinline__ float w8_silu(float x) { return x / (1.0f + __expf(-x)); }

// four biased bytes of `w` -> two bf16x2 words (exact)
__device__ __forceinline__ void w8_widen4(uint32_t w, uint32_t& lo, uint32_t& hi) {
  const float f0 = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7650)) - 8388736.0f;  // 2^23 + 128
  const float f1 = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7651)) - 8388736.0f;
  const float f2 = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7652)) - 8388736.0f;
  const float f3 = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7653)) - 8388736.0f;
  // integers of at most 8 significant bits: the bf16 is the upper half of the fp32, no rounding -> byte permutes
  lo = __byte_perm(__float_as_uint(f0), __float_as_uint(f1), 0x7632);
  hi = __byte_perm(__float_as_uint(f2), __float_as_uint(f3), 0x7632);
}

__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&v)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
               : "memory");
}

template <int BN, int EPI, int NCW>
__global__ void __launch_bounds__(64 + 32 * NCW, 2)
gemm_w8_kernel(const __grid_constant__ CUtensorMap tma_w, const __grid_constant__ CUtensorMap tma_x, const W8Params p) {
  constexpr int B_TILE_BYTES = BN * BK * 2;
  constexpr int STAGE_BYTES = 2 * B_TILE_BYTES + W8_RAW_BYTES;   // raw stage: two activation tiles (1024-aligned), then int8 weights
  constexpr int ACC_COLS = BN < 32 ? 32 : BN;
  constexpr int TMEM_COLS = 256;   // accumulator + 4 x 32 columns of widened weights, rounded up to a power of two
  constexpr bool SILU = EPI == DLI_EPI_SILU_MUL_BF16;
  constexpr int W8_CONVERTERS = 32 * NCW;
  constexpr int KPARTS = NCW / 4;            // warps that share a TMEM lane quarter split the k-block's 64 columns
  constexpr int CHUNKS = 4 / KPARTS;         // 16-byte chunks (16 int8) per thread and k-block: 2 or 1

  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bars[2 * W8_MAX_STAGES + 2 * W8_A_SLOTS + 1];
  __shared__ uint32_t tmem_base_slot;

  const uint32_t tiles = (smem_u32(smem_raw) + 1023u) & ~1023u;      // the raw ring
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int stages = p.stages;
  const uint32_t raw_full0 = smem_u32(&bars[0]), empty0 = smem_u32(&bars[W8_MAX_STAGES]),
                 a_full0 = smem_u32(&bars[2 * W8_MAX_STAGES]), a_empty0 = smem_u32(&bars[2 * W8_MAX_STAGES + W8_A_SLOTS]),
                 tmem_full = smem_u32(&bars[2 * W8_MAX_STAGES + 2 * W8_A_SLOTS]);

  const int kb0 = blockIdx.z * p.kb_per_split;   // kb_per_split is even: a raw stage is a PAIR of k-blocks
  const int kb1 = min(kb0 + p.kb_per_split, p.k_blocks);
  const int kp0 = kb0 >> 1, kp1 = (kb1 + 1) >> 1;   // an odd tail reads zeros (TMA out-of-bounds fill) against zero activations
  const int a_row0 = blockIdx.x * (SILU ? BM / 2 : BM);
  const int b_row0 = blockIdx.y * BN;

  pdl_trigger();
  if (threadIdx.x == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_w) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_x) : "memory");
    for (int s = 0; s < stages; ++s) {
      mbar_init(raw_full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, 1);
    }
    for (int c = 0; c < W8_A_SLOTS; ++c) {
      mbar_init(a_full0 + 8 * c, NCW);   // one arrival per converter warp
      mbar_init(a_empty0 + 8 * c, 1);
    }
    mbar_init(tmem_full, 1);
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

  if (kb1 > kb0) {
    if (warp == 0) {
      // ===== TMA producer: int8 weight tiles (before the dependency wait: weights never depend on the predecessor)
      //       and the activation tiles (after it) =====
      auto load_weights = [&](int s, int kp) {
        const uint32_t raw = tiles + s * STAGE_BYTES + 2 * B_TILE_BYTES;
        if (SILU) {  // 64 gate rows + 64 up rows of the same 64 features
          tma_load_2d(raw, &tma_w, raw_full0 + 8 * s, kp * (W8_RAW_K / W8_TMAP_ELT), a_row0);
          tma_load_2d(raw + W8_RAW_BYTES / 2, &tma_w, raw_full0 + 8 * s, kp * (W8_RAW_K / W8_TMAP_ELT), p.up_row_offset + a_row0);
        } else {
          tma_load_2d(raw, &tma_w, raw_full0 + 8 * s, kp * (W8_RAW_K / W8_TMAP_ELT), a_row0);
        }
      };
      auto load_acts = [&](int s, int kp) {
        tma_load_2d(tiles + s * STAGE_BYTES, &tma_x, raw_full0 + 8 * s, (2 * kp) * BK, b_row0);
        tma_load_2d(tiles + s * STAGE_BYTES + B_TILE_BYTES, &tma_x, raw_full0 + 8 * s, (2 * kp + 1) * BK, b_row0);
      };
      const int n_pre = min(stages, kp1 - kp0);
      if (elect_one()) {
        for (int i = 0; i < n_pre; ++i) {
          mbar_expect_tx(raw_full0 + 8 * i, W8_RAW_BYTES + 2 * B_TILE_BYTES);
          load_weights(i, kp0 + i);
        }
      }
      __syncwarp();
      pdl_wait();
      if (elect_one()) {
        for (int i = 0; i < n_pre; ++i) load_acts(i, kp0 + i);
      }
      __syncwarp();
      int s = n_pre == stages ? 0 : n_pre;
      uint32_t ph = n_pre == stages ? 1 : 0;
      for (int kp = kp0 + n_pre; kp < kp1; ++kp) {
        mbar_wait(empty0 + 8 * s, ph ^ 1);
        if (elect_one()) {
          mbar_expect_tx(raw_full0 + 8 * s, W8_RAW_BYTES + 2 * B_TILE_BYTES);
          load_weights(s, kp);
          load_acts(s, kp);
        }
        __syncwarp();
        if (++s == stages) { s = 0; ph ^= 1; }
      }
    } else if (warp == 1) {
      // ===== MMA issuer =====
      constexpr uint32_t idesc = umma_idesc_bf16(BN);
      int s = 0, c = 0;
      uint32_t cph = 0;
      for (int kp = kp0; kp < kp1; ++kp) {
#pragma unroll
        for (int j = 0; j < 2; ++j) {        // the two k-blocks of the raw stage
          mbar_wait(a_full0 + 8 * c, cph);   // the converters passed raw_full[s] before arriving: the activation tiles are there too
          tcgen05_fence_after();
          if (elect_one()) {
            const uint32_t tmem_a = tmem_d + ACC_COLS + c * W8_A_COLS;
            const uint64_t bdesc = umma_desc_sw128(tiles + s * STAGE_BYTES + j * B_TILE_BYTES);
#ifndef W8_EXP_STREAM   // experiment build: no MMAs, no widening -- the TMA / barrier chain alone
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k)
              w8_umma_ts(tmem_d, tmem_a + 8 * k, bdesc + 2 * k, idesc, (kp > kp0 || j > 0 || k > 0) ? 1u : 0u);
#endif
            umma_commit(a_empty0 + 8 * c);   // the TMEM slot may be overwritten
            if (j == 1) {
              umma_commit(empty0 + 8 * s);   // the raw stage (its activation tiles) may be refilled
              if (kp + 1 == kp1) umma_commit(tmem_full);
            }
          }
          __syncwarp();
          if (++c == W8_A_SLOTS) { c = 0; cph ^= 1; }
        }
        if (++s == stages) s = 0;
      }
    } else {
      // ===== converters (main loop), then epilogue =====
      const int e = threadIdx.x - 64;  // 0..255
      const int q = warp & 3;                // TMEM lane quarter this warp may access: weight rows 32 q .. 32 q + 31
      const int half = (warp - 2) >> 2;      // the KPARTS warps of a quarter split the k-block: 64 / KPARTS int8 each
      {
        // this thread's row of the raw tile ([128 rows][128 B], 128-byte TMA swizzle: 16-byte chunk ^= row & 7): of each of
        // the stage's two k-blocks j it widens chunks 4 j + CHUNKS half .. + CHUNKS - 1
        const int r = q * 32 + lane;
        uint32_t src[2][CHUNKS];
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
          for (int i = 0; i < CHUNKS; ++i)
            src[j][i] = 2 * B_TILE_BYTES + r * W8_RAW_K + (((4 * j + CHUNKS * half + i) ^ (r & 7)) << 4);
        const uint32_t t_lane = tmem_d + ((uint32_t)(q * 32) << 16) + ACC_COLS + half * (8 * CHUNKS);
        int s = 0, c = 0;
        uint32_t ph = 0, cph = 0;
#ifdef W8_TIMING
        long long t_raw = 0, t_slot = 0, t_conv = 0, t_sync = 0, t0 = clock64(), t1;
#define W8_T(acc) do { t1 = clock64(); acc += t1 - t0; t0 = t1; } while (0)
#else
#define W8_T(acc) ((void)0)
#endif
        for (int kp = kp0; kp < kp1; ++kp) {
          mbar_wait(raw_full0 + 8 * s, ph, 40);   // back off between polls: 16 spinning converter warps per SM take issue slots from the TMA and MMA warps
          W8_T(t_raw);
          const uint32_t raw_stage = tiles + s * STAGE_BYTES;
          uint32_t w[2][4 * CHUNKS];
#pragma unroll
          for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int i = 0; i < CHUNKS; ++i)
              asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(w[j][4 * i]), "=r"(w[j][4 * i + 1]), "=r"(w[j][4 * i + 2]), "=r"(w[j][4 * i + 3]) : "r"(raw_stage + src[j][i]));
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            uint32_t o[8 * CHUNKS];
#pragma unroll
            for (int i = 0; i < 4 * CHUNKS; ++i) w8_widen4(w[j][i], o[2 * i], o[2 * i + 1]);
            W8_T(t_conv);
            mbar_wait(a_empty0 + 8 * c, cph ^ 1, 20);   // the MMAs that last read this TMEM slot have completed
            tcgen05_fence_after();
            W8_T(t_slot);
#ifndef W8_EXP_STREAM
            if constexpr (CHUNKS == 2) tmem_st16(t_lane + c * W8_A_COLS, o);
            else tmem_st8(t_lane + c * W8_A_COLS, o);
            tmem_st_wait();
#endif
            W8_T(t_conv);
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(a_full0 + 8 * c);    // the slot is full when all eight warps have arrived
            W8_T(t_sync);
            if (++c == W8_A_SLOTS) { c = 0; cph ^= 1; }
          }
          if (++s == stages) { s = 0; ph ^= 1; }
        }
#ifdef W8_TIMING
        if (p.dbg && blockIdx.x == 7 && blockIdx.y == 0 && blockIdx.z == 0 && (e == 0 || e == W8_CONVERTERS - 32)) {
          long long* d = p.dbg + (e ? 8 : 0);
          d[0] = t_raw; d[1] = t_slot; d[2] = t_conv; d[3] = t_sync; d[4] = kb1 - kb0;   // per k-block figures = sums / this
        }
#endif
      }
      pdl_wait();  // the output buffers belong to the predecessor until it has completed
      mbar_wait(tmem_full, 0, 50);
      tcgen05_fence_after();
      const uint32_t taddr = tmem_d + ((uint32_t)(q * 32) << 16);   // the two warps of a quarter split the accumulator columns
      uint32_t v[16];
      // lane = feature, column = token: transposed through the (now idle) operand ring so that global traffic is
      // 128-bit and coalesced along the feature axis
      float* stage_f = reinterpret_cast<float*>(smem_raw + (tiles - smem_u32(smem_raw)));  // [BN tokens][128 lanes] over the idle raw ring
      const int lane_row = q * 32 + lane;
#pragma unroll 1
      for (int c0 = half * 16; c0 < BN; c0 += 16 * KPARTS) {
        tmem_ld16(taddr + c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) stage_f[(c0 + j) * BM + lane_row] = __uint_as_float(v[j]);
      }
      asm volatile("bar.sync 1, %0;" ::"n"(W8_CONVERTERS) : "memory");
      if (SILU) {
        for (int idx = e; idx < BN * 16; idx += W8_CONVERTERS) {
          const int t = idx >> 4, f = (idx & 15) * 4;
          const int tok = b_row0 + t, feat = a_row0 + f;
          if (tok < p.m_tokens && feat < p.n_features) {
            const float4 g4 = *reinterpret_cast<const float4*>(stage_f + t * BM + f);
            const float4 u4 = *reinterpret_cast<const float4*>(stage_f + t * BM + 64 + f);
            float gs[4], us[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const bool ok = feat + j < p.n_features;
              gs[j] = ok ? p.scale[feat + j] : 0.f;
              us[j] = ok ? p.scale[p.up_row_offset + feat + j] : 0.f;
            }
            const float gg[4] = {g4.x * gs[0], g4.y * gs[1], g4.z * gs[2], g4.w * gs[3]};
            const float uu[4] = {u4.x * us[0], u4.y * us[1], u4.z * us[2], u4.w * us[3]};
            __nv_bfloat16* o = (__nv_bfloat16*)p.out + (long long)tok * p.ldo + feat;
            if (feat + 4 <= p.n_features) {
              *reinterpret_cast<uint2*>(o) = make_uint2(pack_bf16(w8_silu(gg[0]) * uu[0], w8_silu(gg[1]) * uu[1]),
                                                        pack_bf16(w8_silu(gg[2]) * uu[2], w8_silu(gg[3]) * uu[3]));
            } else {
              for (int j = 0; j < 4 && feat + j < p.n_features; ++j) o[j] = __float2bfloat16_rn(w8_silu(gg[j]) * uu[j]);
            }
          }
        }
      } else {
        for (int idx = e; idx < BN * 32; idx += W8_CONVERTERS) {
          const int t = idx >> 5, f = (idx & 31) * 4;
          const int tok = b_row0 + t, feat = a_row0 + f;
          if (tok < p.m_tokens && feat < p.n_features) {
            const float4 a4 = *reinterpret_cast<const float4*>(stage_f + t * BM + f);
            float aa[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) aa[j] *= (feat + j < p.n_features) ? p.scale[feat + j] : 0.f;
            const bool full = feat + 4 <= p.n_features;
            if (EPI == DLI_EPI_STORE_BF16) {
              __nv_bfloat16* o = (__nv_bfloat16*)p.out + (long long)tok * p.ldo + feat;
              if (full) *reinterpret_cast<uint2*>(o) = make_uint2(pack_bf16(aa[0], aa[1]), pack_bf16(aa[2], aa[3]));
              else for (int j = 0; j < 4 && feat + j < p.n_features; ++j) o[j] = __float2bfloat16_rn(aa[j]);
            } else {
              float* o = (float*)p.out + (long long)tok * p.ldo + feat;
              if (full && EPI == DLI_EPI_ADD_F32) {
                asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(o), "f"(aa[0]), "f"(aa[1]), "f"(aa[2]), "f"(aa[3])
                             : "memory");
              } else if (full) {
                *reinterpret_cast<float4*>(o) = make_float4(aa[0], aa[1], aa[2], aa[3]);
              } else {
                for (int j = 0; j < 4 && feat + j < p.n_features; ++j) {
                  if (EPI == DLI_EPI_ADD_F32) atomicAdd(o + j, aa[j]);
                  else o[j] = aa[j];
                }
              }
            }
          }
        }
      }
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"((uint32_t)TMEM_COLS) : "memory");
  }
}

// ---- host side ----------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFnW8)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// 2-D uint8 tensor [rows, cols]; box = [box_rows, 128 bytes], 128-byte swizzle (the converter warps read one row per lane)
int make_tmap_u8(CUtensorMap* map, const void* base, long long rows, long long cols, int box_rows) {
  static EncodeTiledFnW8 fn = nullptr;
  if (!fn) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFnW8>(ptr);
  }
  if (!fn) return fail(DLI_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (cols % 16) != 0)
    return fail(DLI_ERR_ARG, "int8 TMA operand needs a 16-byte aligned base and a row length that is a multiple of 16");
  cuuint64_t dims[2] = {(cuuint64_t)cols / W8_TMAP_ELT, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cols};
  cuuint32_t box[2] = {(cuuint32_t)W8_RAW_K / W8_TMAP_ELT, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, W8_TMAP_ELT == 4 ? CU_TENSOR_MAP_DATA_TYPE_UINT32 : CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(DLI_ERR_CUDA, "cuTensorMapEncodeTiled (uint8) failed with CUresult %d", (int)r);
  return DLI_OK;
}

template <int BN, int EPI, int NCW>
static int launch_w8_variant(const CUtensorMap& tw, const CUtensorMap& tx, const W8Params& p, dim3 grid, cudaStream_t st) {
  constexpr int STAGE_BYTES = 2 * BN * BK * 2 + W8_RAW_BYTES;
  const size_t smem = (size_t)p.stages * STAGE_BYTES + 1024;
  auto kern = gemm_w8_kernel<BN, EPI, NCW>;
  DLI_ONCE_PER_DEVICE(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 1024));
  return launch_kernel("gemm_w8_kernel", kern, grid, dim3(64 + 32 * NCW), smem, st, tw, tx, p);
}

template <int BN>
static int launch_w8_epi(int epi, const CUtensorMap& tw, const CUtensorMap& tx, const W8Params& p, dim3 grid, cudaStream_t st) {
  switch (epi) {
    case DLI_EPI_STORE_F32: return launch_w8_variant<BN, DLI_EPI_STORE_F32, 8>(tw, tx, p, grid, st);
    case DLI_EPI_ADD_F32: return launch_w8_variant<BN, DLI_EPI_ADD_F32, 8>(tw, tx, p, grid, st);
    case DLI_EPI_STORE_BF16: return launch_w8_variant<BN, DLI_EPI_STORE_BF16, 8>(tw, tx, p, grid, st);
    case DLI_EPI_SILU_MUL_BF16: return launch_w8_variant<BN, DLI_EPI_SILU_MUL_BF16, 8>(tw, tx, p, grid, st);   // 16 warps measured no faster (the CTAs wait for data)
    default: return fail(DLI_ERR_ARG, "unknown epilogue %d", epi);
  }
}

// tmap_w: uint8 map of the quantised matrix with box rows = 128 (64 for the gate|up epilogue); tmap_x: bf16 activations
// with box rows = plan.bn.  The plan is plan_gemm's (same tiles and split-K); only the stage count differs.
int launch_gemm_w8_planned(const GemmPlan& g, const CUtensorMap& tmap_x, const CUtensorMap& tmap_w, const float* scale, void* out,
                           long long ldo, int m, int n, int epilogue, cudaStream_t st) {
  if (!g.swap) return fail(DLI_ERR_UNSUPPORTED, "the int8-weight GEMM is the decode kernel: m = %d > 128", m);
  W8Params p{};
  p.out = out;
  p.ldo = ldo;
  p.m_tokens = m;
  p.n_features = n;
  p.k_blocks = g.k_blocks;
  p.kb_per_split = (g.kb_per_split + 1) & ~1;   // whole pairs of k-blocks per split
  const int splits = (g.k_blocks + p.kb_per_split - 1) / p.kb_per_split;
  const int stage_bytes = 2 * g.bn * BK * 2 + W8_RAW_BYTES;
  // two CTAs per SM: raw ring <= ~111 KB; the epilogue's transposition buffer (bn x 512 B) overlays it
  int stages = std::min(W8_MAX_STAGES, (112 * 1024 - 2048) / stage_bytes);
  while (stages * stage_bytes < g.bn * 512) ++stages;
  p.stages = std::max(2, stages);
  p.up_row_offset = n;
  p.scale = scale;
  p.dbg = nullptr;
#ifdef W8_TIMING
  if (const char* e = getenv("DLI_W8_DBG_PTR")) p.dbg = reinterpret_cast<long long*>(strtoull(e, nullptr, 16));
#endif
  dim3 grid(g.grid_x, g.grid_y, splits);
  switch (g.bn) {
    case 16: return launch_w8_epi<16>(epilogue, tmap_w, tmap_x, p, grid, st);
    case 32: return launch_w8_epi<32>(epilogue, tmap_w, tmap_x, p, grid, st);
    case 64: return launch_w8_epi<64>(epilogue, tmap_w, tmap_x, p, grid, st);
    default: return launch_w8_epi<128>(epilogue, tmap_w, tmap_x, p, grid, st);
  }
}

int launch_gemm_w8(const void* x, long long ldx, const void* q, const float* scale, void* out, long long ldo, int m, int n,
                   int k, int epilogue, cudaStream_t st) {
  DLI_REQUIRE(x && q && scale && out, "dli_gemm_w8: null pointer");
  DLI_REQUIRE(m > 0 && m <= 128 && n > 0 && k > 0, "dli_gemm_w8: m=%d n=%d k=%d (decode kernel: m <= 128)", m, n, k);
  DLI_REQUIRE(k % 16 == 0 && ldx % 8 == 0, "dli_gemm_w8: k must be a multiple of 16 and ldx of 8 (k=%d ldx=%lld)", k, ldx);
  DLI_REQUIRE(epilogue >= 0 && epilogue <= DLI_EPI_SILU_MUL_BF16, "dli_gemm_w8: unknown epilogue %d", epilogue);
  int epi = epilogue;
  GemmPlan g = plan_gemm(m, n, k, epi);
  if (epi == DLI_EPI_STORE_F32) {  // split-K partials meet in L2: clear the destination, then add
    DLI_CUDA_OK(cudaMemset2DAsync(out, (size_t)ldo * 4, 0, (size_t)n * 4, (size_t)m, st));
    count_launch();
    epi = DLI_EPI_ADD_F32;
    g = plan_gemm(m, n, k, epi);
  }
  const long long w_rows = epilogue == DLI_EPI_SILU_MUL_BF16 ? 2LL * n : n;
  CUtensorMap tx, tw;
  int rc = make_tmap_bf16(&tx, x, m, k, ldx, g.b_box_rows);
  if (rc) return rc;
  if ((rc = make_tmap_u8(&tw, q, w_rows, k, g.a_box_rows))) return rc;
  return launch_gemm_w8_planned(g, tx, tw, scale, out, ldo, m, n, epi, st);
}

// ---------------------------------------------------------------------------------------------------
// Row-wise absmax int8 quantisation (the weight format of bitsandbytes Linear8bitLt with has_fp16_weights=False,
// utils/model.py:102-108): scale[r] = max_k |w[r, k]| / 127, q[r, k] = round(w[r, k] / scale[r]) stored as q + 128.
// One CTA per row; and the inverse, for the prefill path, which runs the bf16 tensor-core GEMMs on a dequantised copy.
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) quantize_rows_w8_kernel(const __nv_bfloat16* __restrict__ w, long long k,
                                                               uint8_t* __restrict__ q, float* __restrict__ scale) {
  __shared__ float red[8];
  const long long row = blockIdx.x;
  const __nv_bfloat16* src = w + row * k;
  float mx = 0.f;
  for (long long i = threadIdx.x * 8LL; i < k; i += 256 * 8) {
    const uint4 a = *reinterpret_cast<const uint4*>(src + i);
    const uint32_t u[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) mx = fmaxf(mx, fmaxf(fabsf(bf16_lo(u[j])), fabsf(bf16_hi(u[j]))));
  }
  mx = warp_max(mx);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
  __syncthreads();
  mx = red[0];
#pragma unroll
  for (int i = 1; i < 8; ++i) mx = fmaxf(mx, red[i]);
  const float s = mx / 127.0f, inv = mx > 0.f ? 127.0f / mx : 0.f;
  if (threadIdx.x == 0) scale[row] = s;
  uint8_t* dst = q + row * k;
  for (long long i = threadIdx.x * 8LL; i < k; i += 256 * 8) {
    const uint4 a = *reinterpret_cast<const uint4*>(src + i);
    const uint32_t u[4] = {a.x, a.y, a.z, a.w};
    uint32_t o[2] = {0, 0};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int q0 = max(-127, min(127, __float2int_rn(bf16_lo(u[j]) * inv))) + 128;
      const int q1 = max(-127, min(127, __float2int_rn(bf16_hi(u[j]) * inv))) + 128;
      o[j >> 1] |= ((uint32_t)q0 | ((uint32_t)q1 << 8)) << (16 * (j & 1));
    }
    *reinterpret_cast<uint2*>(dst + i) = make_uint2(o[0], o[1]);
  }
}

__global__ void __launch_bounds__(256) dequantize_rows_w8_kernel(const uint8_t* __restrict__ q, const float* __restrict__ scale,
                                                                 long long k, __nv_bfloat16* __restrict__ w) {
  pdl_trigger();
  pdl_wait();
  const long long row = blockIdx.x;
  const float s = scale[row];
  const uint8_t* src = q + row * k;
  __nv_bfloat16* dst = w + row * k;
  for (long long i = (blockIdx.y * 256LL + threadIdx.x) * 8; i < k; i += (long long)gridDim.y * 256 * 8) {
    const uint2 a = *reinterpret_cast<const uint2*>(src + i);
    const uint32_t u[2] = {a.x, a.y};
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t b = u[j >> 1] >> (16 * (j & 1));
      o[j] = pack_bf16(((int)(b & 0xff) - 128) * s, ((int)((b >> 8) & 0xff) - 128) * s);
    }
    *reinterpret_cast<uint4*>(dst + i) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

int launch_quantize_rows_w8(const void* w_bf16, long long n, long long k, void* q_u8, float* scale, cudaStream_t st) {
  DLI_REQUIRE(w_bf16 && q_u8 && scale && n > 0 && k > 0 && k % 16 == 0, "dli_quantize_rows_w8: null pointer or k %% 16 != 0");
  quantize_rows_w8_kernel<<<(unsigned)n, 256, 0, st>>>((const __nv_bfloat16*)w_bf16, k, (uint8_t*)q_u8, scale);
  DLI_CHECK_LAUNCH("quantize_rows_w8_kernel");
  return DLI_OK;
}

int launch_dequantize_rows_w8(const void* q_u8, const float* scale, long long n, long long k, void* w_bf16, cudaStream_t st) {
  DLI_REQUIRE(w_bf16 && q_u8 && scale && n > 0 && k > 0 && k % 16 == 0, "dli_dequantize_rows_w8: null pointer or k %% 16 != 0");
  const int ny = (int)std::max<long long>(1, std::min<long long>(4, k / 2048));
  return launch_kernel("dequantize_rows_w8_kernel", dequantize_rows_w8_kernel, dim3((unsigned)n, ny), dim3(256), 0, st,
                       (const uint8_t*)q_u8, scale, k, (__nv_bfloat16*)w_bf16);
}

}  // namespace dli
