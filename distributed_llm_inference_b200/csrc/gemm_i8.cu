// int8 x int8 decode GEMM (LLM.int8 arithmetic):
//   out[m, n] (+)= w_scale[n] * ( s1[m] * sum_k c1[m, k] * (q[n, k] - 128)  +  s2[m] * sum_k c2[m, k] * (q[n, k] - 128) )
//
// The reference's optimised mode (utils/model.py:93-123) is bitsandbytes Linear8bitLt(threshold=5.0,
// has_fp16_weights=False): weights AND activations in int8 (row-wise absmax scales on both sides), multiplied in int32,
// with the activations whose magnitude reaches the threshold ("outliers") kept out of that product so that they do not
// inflate the row's scale.  gemm_w8.cu keeps the activations in bf16 and pays for widening the weights in registers; here
// both operands go to the tensor core as they are: tcgen05.mma.kind::i8 (A = biased uint8 weights, B = signed int8
// activations, int32 accumulators in TMEM), so the kernel is the bf16 weight-streaming kernel of gemm_tcgen05.cu byte for
// byte -- TMA lands 128-byte K slices (now 128 elements) with the 128B swizzle, four MMAs of K = 32 per k-block, swap-AB,
// split-K over gridDim.z -- at half the weight bytes and with no conversion work.
//   * Outliers are a SECOND int8 plane of the activation (c2: the entries at or above the threshold, quantised with the
//     row's absmax over those entries; c1 holds 0 there and vice versa): every weight tile in shared memory is
//     multiplied by both planes at once (their tiles are adjacent rows of the B operand: one MMA of N = 2 BN).  bitsandbytes multiplies its outlier columns in fp16; a second
//     8-bit plane keeps them to 1/254 of the row's largest outlier -- the resolution the regular entries have relative
//     to the threshold -- at a cost that does not depend on how many there are (a first cut added them in the epilogue
//     from per-row lists, reading weight columns byte by byte: with ~100 scattered outliers per row, as silu(gate) * up
//     of a random-init model has, one GEMM took 175 us longer).
//   * The weights are stored biased (q + 128, the format of dli_quantize_rows_w8): sum_k c * q_u = D, and
//     sum_k c * (q_u - 128) = D - 128 * sum_k c, exact in int32; the activation quantiser leaves sum_k c per k-block and
//     plane (ksum[m][kb]) so that a split-K CTA subtracts the sum over its own k range.
//   * Epilogue: int32 -> fp32 times the token's plane scales, planes added, transposed through shared memory, times the
//     feature's weight scale, then the usual forms (fp32 store / red.add into the residual / silu(gate) * up in bf16).
#include "common.cuh"
#include "tc_ptx.cuh"

#include <algorithm>
#include <stdlib.h>

namespace dli {

constexpr int I8_BK = 128;                       // elements (= bytes) per k-block
constexpr int I8_A_TILE_BYTES = BM * I8_BK;      // 128 weight rows x 128 B
constexpr int I8_THREADS = 192;                  // TMA warp, MMA warp, four epilogue warps
constexpr int I8_MAX_STAGES = 8;
constexpr int I8_UMMA_K = 32;

struct I8Params {
  void* out;
  long long ldo;
  int m_tokens, n_features;
  int k_blocks, kb_per_split;
  int stages;
  int up_row_offset;          // SILU epilogue: row of q / w_scale where up_proj starts (= n_features)
  const float* w_scale;       // [rows of q]
  const float* x_scale[2];    // [m] per plane
  const int32_t* x_sum[2];    // [m][k_blocks] per plane
};

// kind::i8 instruction descriptor: D = s32, A = unsigned 8-bit (biased weights), B = signed 8-bit, both K-major, M = 128
__host__ __device__ constexpr uint32_t umma_idesc_i8(int bn) {
  return (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t)(bn >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
}
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ float i8_silu(float x) { return x / (1.0f + __expf(-x)); }

template <int BN, int EPI, int PLANES>
__global__ void __launch_bounds__(I8_THREADS, 2)
gemm_i8_kernel(const __grid_constant__ CUtensorMap tma_w, const __grid_constant__ CUtensorMap tma_x, const __grid_constant__ CUtensorMap tma_x2,
               const I8Params p) {
  constexpr int B_TILE_BYTES = BN * I8_BK;
  constexpr int STAGE_BYTES = I8_A_TILE_BYTES + PLANES * B_TILE_BYTES;
  constexpr int MMA_N = PLANES * BN;              // both planes are ONE MMA: their tiles are adjacent rows of the B operand
  constexpr int TMEM_COLS = MMA_N < 32 ? 32 : MMA_N;   // accumulator columns [0, BN) = first plane, [BN, 2 BN) = outlier plane
  constexpr bool SILU = EPI == DLI_EPI_SILU_MUL_BF16;

  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bars[2 * I8_MAX_STAGES + 1];
  __shared__ uint32_t tmem_base_slot;
  __shared__ int xs_s[PLANES][BN];     // 128 * (sum of the token's codes over this CTA's k range), per plane
  __shared__ float sx_s[PLANES][BN];

  const uint32_t tiles = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int stages = p.stages;
  const uint32_t full0 = smem_u32(&bars[0]), empty0 = smem_u32(&bars[I8_MAX_STAGES]);
  const uint32_t tmem_full = smem_u32(&bars[2 * I8_MAX_STAGES]);

  const int kb0 = blockIdx.z * p.kb_per_split;
  const int kb1 = min(kb0 + p.kb_per_split, p.k_blocks);
  const int a_row0 = blockIdx.x * (SILU ? BM / 2 : BM);
  const int b_row0 = blockIdx.y * BN;

  pdl_trigger();
  if (threadIdx.x == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_w) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_x) : "memory");
    if (PLANES == 2) asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_x2) : "memory");
    for (int s = 0; s < stages; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, 1);
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
      // ===== TMA producer: weight tiles before the dependency wait (they never depend on the predecessor), the
      //       quantised activations after it =====
      auto load_weights = [&](int s, int kb) {
        const uint32_t a_dst = tiles + s * STAGE_BYTES;
        if (SILU) {  // 64 gate rows + 64 up rows of the same 64 features
          tma_load_2d(a_dst, &tma_w, full0 + 8 * s, kb * I8_BK, a_row0);
          tma_load_2d(a_dst + I8_A_TILE_BYTES / 2, &tma_w, full0 + 8 * s, kb * I8_BK, p.up_row_offset + a_row0);
        } else {
          tma_load_2d(a_dst, &tma_w, full0 + 8 * s, kb * I8_BK, a_row0);
        }
      };
      auto load_acts = [&](int s, int kb) {
        const uint32_t b_dst = tiles + s * STAGE_BYTES + I8_A_TILE_BYTES;
        tma_load_2d(b_dst, &tma_x, full0 + 8 * s, kb * I8_BK, b_row0);
        if (PLANES == 2) tma_load_2d(b_dst + B_TILE_BYTES, &tma_x2, full0 + 8 * s, kb * I8_BK, b_row0);
      };
      const int n_pre = min(stages, kb1 - kb0);
      if (elect_one()) {
        for (int i = 0; i < n_pre; ++i) {
          mbar_expect_tx(full0 + 8 * i, STAGE_BYTES);
          load_weights(i, kb0 + i);
        }
      }
      __syncwarp();
      pdl_wait();
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
    } else if (warp == 1) {
      // ===== MMA issuer: every weight tile against both activation planes (one MMA of N = 2 BN) =====
      constexpr uint32_t idesc = umma_idesc_i8(MMA_N);
      int s = 0;
      uint32_t ph = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(full0 + 8 * s, ph);
        tcgen05_fence_after();
        if (elect_one()) {
          const uint32_t a_src = tiles + s * STAGE_BYTES;
          const uint64_t adesc = umma_desc_sw128(a_src), bdesc = umma_desc_sw128(a_src + I8_A_TILE_BYTES);
#pragma unroll
          for (int k = 0; k < I8_BK / I8_UMMA_K; ++k)   // 32 int8 = 32 bytes along K inside the swizzle atom: +2 in the (addr >> 4) field
            umma_i8(tmem_d, adesc + 2 * k, bdesc + 2 * k, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          umma_commit(empty0 + 8 * s);
          if (kb + 1 == kb1) umma_commit(tmem_full);
        }
        __syncwarp();
        if (++s == stages) { s = 0; ph ^= 1; }
      }
    } else {
      // ===== epilogue: int32 accumulators -> token scales -> transposition -> weight scales -> global =====
      pdl_wait();  // the output buffers and the activation's scales / sums belong to the predecessor until it has completed
      const int e = threadIdx.x - 64;  // 0..127
      if (e < BN) {
        const int tok = b_row0 + e;
#pragma unroll
        for (int pl = 0; pl < PLANES; ++pl) {
          int s = 0;
          float sx = 0.f;
          if (tok < p.m_tokens) {
            const int32_t* xs = p.x_sum[pl] + (long long)tok * p.k_blocks;
            for (int kb = kb0; kb < kb1; ++kb) s += xs[kb];
            sx = p.x_scale[pl][tok];
          }
          xs_s[pl][e] = 128 * s;
          sx_s[pl][e] = sx;
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");  // the four epilogue warps only: xs_s / sx_s are published
      mbar_wait(tmem_full, 0, 200);
      tcgen05_fence_after();
      const int q4 = warp & 3;  // TMEM lane quarter this warp may read
      const uint32_t taddr = tmem_d + ((uint32_t)(q4 * 32) << 16);
      uint32_t v[16], u[16];
      // lane = feature, column = token: transposed through the (now idle) operand ring so that global traffic is 128-bit
      // and coalesced along the feature axis
      float* stage_f = reinterpret_cast<float*>(smem_raw + (tiles - smem_u32(smem_raw)));  // [BN tokens][128 lanes]
      const int lane_row = q4 * 32 + lane;
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 16) {
        tmem_ld16(taddr + c0, v);
        if (PLANES == 2) tmem_ld16(taddr + BN + c0, u);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          float f = (float)((int)v[j] - xs_s[0][c0 + j]) * sx_s[0][c0 + j];
          if (PLANES == 2) f += (float)((int)u[j] - xs_s[PLANES - 1][c0 + j]) * sx_s[PLANES - 1][c0 + j];
          stage_f[(c0 + j) * BM + lane_row] = f;
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (SILU) {
        // lanes 0..63 hold gate, 64..127 up of features a_row0 .. a_row0 + 63; 4 features per thread and token
        for (int idx = e; idx < BN * 16; idx += 128) {
          const int t = idx >> 4, f = (idx & 15) * 4;
          const int tok = b_row0 + t, feat = a_row0 + f;
          if (tok < p.m_tokens && feat + 4 <= p.n_features) {   // n_features % 4 == 0 (checked by the launcher)
            const float4 g = *reinterpret_cast<const float4*>(stage_f + t * BM + f);
            const float4 up = *reinterpret_cast<const float4*>(stage_f + t * BM + 64 + f);
            const float4 sg = *reinterpret_cast<const float4*>(p.w_scale + feat);
            const float4 su = *reinterpret_cast<const float4*>(p.w_scale + p.up_row_offset + feat);
            __nv_bfloat16* o = (__nv_bfloat16*)p.out + (long long)tok * p.ldo + feat;
            *reinterpret_cast<uint2*>(o) =
                make_uint2(pack_bf16(i8_silu(g.x * sg.x) * (up.x * su.x), i8_silu(g.y * sg.y) * (up.y * su.y)),
                           pack_bf16(i8_silu(g.z * sg.z) * (up.z * su.z), i8_silu(g.w * sg.w) * (up.w * su.w)));
          }
        }
      } else {
        for (int idx = e; idx < BN * 32; idx += 128) {
          const int t = idx >> 5, f = (idx & 31) * 4;
          const int tok = b_row0 + t, feat = a_row0 + f;
          if (tok < p.m_tokens && feat + 4 <= p.n_features) {
            float4 a = *reinterpret_cast<const float4*>(stage_f + t * BM + f);
            const float4 sw = *reinterpret_cast<const float4*>(p.w_scale + feat);
            a.x *= sw.x; a.y *= sw.y; a.z *= sw.z; a.w *= sw.w;
            if (EPI == DLI_EPI_STORE_BF16) {
              __nv_bfloat16* o = (__nv_bfloat16*)p.out + (long long)tok * p.ldo + feat;
              *reinterpret_cast<uint2*>(o) = make_uint2(pack_bf16(a.x, a.y), pack_bf16(a.z, a.w));
            } else {
              float* o = (float*)p.out + (long long)tok * p.ldo + feat;
              if (EPI == DLI_EPI_ADD_F32) {
                asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(o), "f"(a.x), "f"(a.y), "f"(a.z), "f"(a.w)
                             : "memory");
              } else {
                *reinterpret_cast<float4*>(o) = a;
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
template <int BN, int EPI, int PLANES>
static int launch_i8_variant(const CUtensorMap& tw, const CUtensorMap& tx, const CUtensorMap& tx2, const I8Params& p, dim3 grid,
                             cudaStream_t st) {
  constexpr int STAGE_BYTES = I8_A_TILE_BYTES + PLANES * BN * I8_BK;
  const size_t smem = (size_t)p.stages * STAGE_BYTES + 1024;
  auto kern = gemm_i8_kernel<BN, EPI, PLANES>;
  DLI_ONCE_PER_DEVICE(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
  return launch_kernel("gemm_i8_kernel", kern, grid, dim3(I8_THREADS), smem, st, tw, tx, tx2, p);
}

template <int BN, int PLANES>
static int launch_i8_epi(int epi, const CUtensorMap& tw, const CUtensorMap& tx, const CUtensorMap& tx2, const I8Params& p, dim3 grid,
                         cudaStream_t st) {
  switch (epi) {
    case DLI_EPI_STORE_F32: return launch_i8_variant<BN, DLI_EPI_STORE_F32, PLANES>(tw, tx, tx2, p, grid, st);
    case DLI_EPI_ADD_F32: return launch_i8_variant<BN, DLI_EPI_ADD_F32, PLANES>(tw, tx, tx2, p, grid, st);
    case DLI_EPI_STORE_BF16: return launch_i8_variant<BN, DLI_EPI_STORE_BF16, PLANES>(tw, tx, tx2, p, grid, st);
    case DLI_EPI_SILU_MUL_BF16: return launch_i8_variant<BN, DLI_EPI_SILU_MUL_BF16, PLANES>(tw, tx, tx2, p, grid, st);
    default: return fail(DLI_ERR_ARG, "unknown epilogue %d", epi);
  }
}

// The byte geometry is the bf16 decode kernel's (128-byte K slices, 16 KB weight tiles), so the plan is plan_gemm's for
// a bf16 problem of k / 2: same tiles and split-K.
GemmPlan plan_gemm_i8(int m, int n, int k, int epilogue) { return plan_gemm(m, n, (k + 1) / 2, epilogue); }

// tmap_w: uint8 map of the quantised matrix with box rows = plan.a_box_rows; tmap_x / tmap_x2: uint8 maps of x.q / x.q2
// with box rows = plan.bn (tmap_x2 is ignored for a one-plane activation)
int launch_gemm_i8_planned(const GemmPlan& g, const CUtensorMap& tmap_x, const CUtensorMap& tmap_x2, const CUtensorMap& tmap_w,
                           const dli_act_i8& x, const float* w_scale, void* out, long long ldo, int m, int n, int epilogue,
                           cudaStream_t st) {
  if (!g.swap) return fail(DLI_ERR_UNSUPPORTED, "the int8 x int8 GEMM is the decode kernel: m = %d > 128", m);
  if (n % 4) return fail(DLI_ERR_ARG, "dli_gemm_i8: n = %d must be a multiple of 4", n);
  if (g.k_blocks > 512) return fail(DLI_ERR_UNSUPPORTED, "dli_gemm_i8: k > 65536 could overflow the int32 accumulators (127 * 255 * k)");
  const int planes = x.q2 ? 2 : 1;
  I8Params p{};
  p.out = out;
  p.ldo = ldo;
  p.m_tokens = m;
  p.n_features = n;
  p.k_blocks = g.k_blocks;
  p.kb_per_split = g.kb_per_split;
  const int stage_bytes = I8_A_TILE_BYTES + planes * g.bn * I8_BK;
  static int ring_kb = 0;   // shared memory per CTA for the operand ring; experiments build: DLI_I8_RING_KB
  if (!ring_kb) {
    const char* e = exp_env("DLI_I8_RING_KB");
    ring_kb = e ? std::max(24, atoi(e)) : 110;
  }
  int stages = std::min(I8_MAX_STAGES, (ring_kb * 1024) / stage_bytes);   // 110 KB: two CTAs per SM
  stages = std::max(2, std::min(stages, g.kb_per_split));
  while (stages * stage_bytes < g.bn * 512) ++stages;   // the epilogue's transposition buffer overlays the ring
  p.stages = stages;
  p.up_row_offset = n;
  p.w_scale = w_scale;
  p.x_scale[0] = x.scale;
  p.x_sum[0] = x.ksum;
  p.x_scale[1] = x.scale2;
  p.x_sum[1] = x.ksum2;
  dim3 grid(g.grid_x, g.grid_y, g.splits);
#define DLI_I8_BN(BN_)                                                                             \
  return planes == 2 ? launch_i8_epi<BN_, 2>(epilogue, tmap_w, tmap_x, tmap_x2, p, grid, st)         \
                     : launch_i8_epi<BN_, 1>(epilogue, tmap_w, tmap_x, tmap_x, p, grid, st)
  switch (g.bn) {
    case 16: DLI_I8_BN(16);
    case 32: DLI_I8_BN(32);
    case 64: DLI_I8_BN(64);
    default: DLI_I8_BN(128);
  }
#undef DLI_I8_BN
}

int launch_gemm_i8(const dli_act_i8* x, const void* q, const float* w_scale, void* out, long long ldo, int m, int n, int k,
                   int epilogue, cudaStream_t st) {
  DLI_REQUIRE(x && x->q && x->scale && x->ksum && q && w_scale && out, "dli_gemm_i8: null pointer");
  DLI_REQUIRE(!x->q2 || (x->scale2 && x->ksum2), "dli_gemm_i8: the outlier plane needs its scales and sums");
  DLI_REQUIRE(m > 0 && m <= 128 && n > 0 && k > 0, "dli_gemm_i8: m=%d n=%d k=%d (decode kernel: m <= 128)", m, n, k);
  DLI_REQUIRE(k % 16 == 0, "dli_gemm_i8: k = %d must be a multiple of 16", k);
  DLI_REQUIRE(epilogue >= 0 && epilogue <= DLI_EPI_SILU_MUL_BF16, "dli_gemm_i8: unknown epilogue %d", epilogue);
  int epi = epilogue;
  GemmPlan g = plan_gemm_i8(m, n, k, epi);
  if (epi == DLI_EPI_STORE_F32) {  // split-K partials meet in L2: clear the destination, then add
    DLI_CUDA_OK(cudaMemset2DAsync(out, (size_t)ldo * 4, 0, (size_t)n * 4, (size_t)m, st));
    count_launch();
    epi = DLI_EPI_ADD_F32;
    g = plan_gemm_i8(m, n, k, epi);
  }
  const long long w_rows = epilogue == DLI_EPI_SILU_MUL_BF16 ? 2LL * n : n;
  CUtensorMap tx, tx2, tw;
  int rc = make_tmap_u8(&tx, x->q, m, k, g.b_box_rows);
  if (rc) return rc;
  tx2 = tx;
  if (x->q2 && (rc = make_tmap_u8(&tx2, x->q2, m, k, g.b_box_rows))) return rc;
  if ((rc = make_tmap_u8(&tw, q, w_rows, k, g.a_box_rows))) return rc;
  return launch_gemm_i8_planned(g, tx, tx2, tw, *x, w_scale, out, ldo, m, n, epi, st);
}

// ---------------------------------------------------------------------------------------------------
// Activation quantiser: one CTA per token row.  Pass 1: absmax over the entries below the threshold and over the entries
// at or above it.  Pass 2: the two planes of codes (round to nearest even, clamped to +-127; an entry is 0 in the plane
// it does not belong to) and the per-128-block sums of the codes.  `x` is read after the dependency wait.
// ---------------------------------------------------------------------------------------------------
constexpr int QA_THREADS = 512;
constexpr int QA_CHUNKS = 8;   // rows of up to 512 * 8 * 8 = 32768 entries are read once and kept in registers
__global__ void __launch_bounds__(QA_THREADS) quantize_act_i8_kernel(const __nv_bfloat16* __restrict__ x, long long ldx, int k, int k_blocks,
                                                                     float threshold, dli_act_i8 d) {
  __shared__ float red[2][QA_THREADS / 32];
  pdl_trigger();
  const int row = blockIdx.x;
  const __nv_bfloat16* src = x + (long long)row * ldx;
  const bool two = d.q2 != nullptr && threshold > 0.f;
  const float thr = two ? threshold : INFINITY;
  const int i0 = threadIdx.x * 8;
  pdl_wait();
  uint4 a[QA_CHUNKS];
#pragma unroll
  for (int c = 0; c < QA_CHUNKS; ++c) {
    const int i = i0 + c * QA_THREADS * 8;
    if (i < k) a[c] = *reinterpret_cast<const uint4*>(src + i);
  }
  float mx = 0.f, mo = 0.f;
#pragma unroll
  for (int c = 0; c < QA_CHUNKS; ++c) {
    if (i0 + c * QA_THREADS * 8 < k) {
      const uint32_t u[4] = {a[c].x, a[c].y, a[c].z, a[c].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float v0 = fabsf(bf16_lo(u[j])), v1 = fabsf(bf16_hi(u[j]));
        if (v0 < thr) mx = fmaxf(mx, v0); else mo = fmaxf(mo, v0);
        if (v1 < thr) mx = fmaxf(mx, v1); else mo = fmaxf(mo, v1);
      }
    }
  }
  mx = warp_max(mx);
  mo = warp_max(mo);
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = mx; red[1][threadIdx.x >> 5] = mo; }
  __syncthreads();
  mx = red[0][0];
  mo = red[1][0];
#pragma unroll
  for (int i = 1; i < QA_THREADS / 32; ++i) { mx = fmaxf(mx, red[0][i]); mo = fmaxf(mo, red[1][i]); }
  const float inv = mx > 0.f ? 127.0f / mx : 0.f, inv2 = mo > 0.f ? 127.0f / mo : 0.f;
  if (threadIdx.x == 0) {
    d.scale[row] = mx / 127.0f;
    if (d.q2) d.scale2[row] = mo / 127.0f;
  }
  int8_t* dst = (int8_t*)d.q + (long long)row * k;
  int8_t* dst2 = (int8_t*)d.q2 + (long long)row * k;
  int32_t* ks = d.ksum + (long long)row * k_blocks;
  int32_t* ks2 = d.ksum2 + (long long)row * k_blocks;
#pragma unroll
  for (int c = 0; c < QA_CHUNKS; ++c) {   // 16 consecutive threads share a 128-entry block; every lane takes part in the shuffles
    const int i = i0 + c * QA_THREADS * 8;
    if (c * QA_THREADS * 8 >= k) break;   // uniform over the CTA
    int sum = 0, sum2 = 0;
    if (i < k) {
      const float v[8] = {bf16_lo(a[c].x), bf16_hi(a[c].x), bf16_lo(a[c].y), bf16_hi(a[c].y),
                          bf16_lo(a[c].z), bf16_hi(a[c].z), bf16_lo(a[c].w), bf16_hi(a[c].w)};
      uint2 p1, p2;
      i8_codes8(v, thr, inv, inv2, p1, p2, sum, sum2);
      *reinterpret_cast<uint2*>(dst + i) = p1;
      if (d.q2) *reinterpret_cast<uint2*>(dst2 + i) = p2;
    }
#pragma unroll
    for (int off = 8; off > 0; off >>= 1) {
      sum += __shfl_xor_sync(0xffffffffu, sum, off);
      sum2 += __shfl_xor_sync(0xffffffffu, sum2, off);
    }
    if ((threadIdx.x & 15) == 0 && i < k) {
      ks[i >> 7] = sum;
      if (d.q2) ks2[i >> 7] = sum2;
    }
  }
}

int launch_quantize_act_i8(const void* x_bf16, long long ldx, int m, int k, float threshold, const dli_act_i8* dst, cudaStream_t st) {
  DLI_REQUIRE(x_bf16 && dst && dst->q && dst->scale && dst->ksum, "dli_quantize_act_i8: null pointer");
  DLI_REQUIRE(!dst->q2 || (dst->scale2 && dst->ksum2), "dli_quantize_act_i8: the outlier plane needs its scales and sums");
  DLI_REQUIRE(m > 0 && k > 0 && k % 16 == 0 && ldx % 8 == 0, "dli_quantize_act_i8: m=%d k=%d ldx=%lld (k %% 16 == 0, ldx %% 8 == 0)", m, k, ldx);
  DLI_REQUIRE(k <= QA_THREADS * 8 * QA_CHUNKS, "dli_quantize_act_i8: rows of more than %d entries are not supported (k = %d)", QA_THREADS * 8 * QA_CHUNKS, k);
  return launch_kernel("quantize_act_i8_kernel", quantize_act_i8_kernel, dim3((unsigned)m), dim3(QA_THREADS), 0, st,
                       (const __nv_bfloat16*)x_bf16, ldx, k, (k + I8_BK - 1) / I8_BK, threshold, *dst);
}

static size_t i8_align(size_t v) { return (v + 255) & ~(size_t)255; }
size_t act_i8_bytes(int m, int k, int planes) {
  const size_t kb = (size_t)(k + I8_BK - 1) / I8_BK;
  return (size_t)planes * (i8_align((size_t)m * k) + i8_align((size_t)m * 4) + i8_align((size_t)m * kb * 4));
}
int act_i8_carve(void* base, size_t bytes, int m, int k, int planes, dli_act_i8* out) {
  DLI_REQUIRE(base && out && m > 0 && k > 0 && (planes == 1 || planes == 2), "dli_act_i8_carve: bad argument");
  DLI_REQUIRE((reinterpret_cast<uintptr_t>(base) & 255) == 0, "dli_act_i8_carve: base must be 256-byte aligned");
  DLI_REQUIRE(bytes >= act_i8_bytes(m, k, planes), "dli_act_i8_carve: %zu bytes given, %zu needed", bytes, act_i8_bytes(m, k, planes));
  const size_t kb = (size_t)(k + I8_BK - 1) / I8_BK;
  uint8_t* p = (uint8_t*)base;
  *out = dli_act_i8{};
  out->q = p; p += i8_align((size_t)m * k);
  out->scale = (float*)p; p += i8_align((size_t)m * 4);
  out->ksum = (int32_t*)p; p += i8_align((size_t)m * kb * 4);
  if (planes == 2) {
    out->q2 = p; p += i8_align((size_t)m * k);
    out->scale2 = (float*)p; p += i8_align((size_t)m * 4);
    out->ksum2 = (int32_t*)p;
  }
  return DLI_OK;
}

}  // namespace dli
