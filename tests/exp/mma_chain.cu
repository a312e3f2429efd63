// Experiment (not product code): issue rate / latency of tcgen05.mma chains with operands resident in smem.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tests/exp/mma_chain tests/exp/mma_chain.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t desc_sw128(uint32_t a) {
  return (uint64_t)((a >> 4) & 0x3FFF) | ((uint64_t)1 << 16) | ((uint64_t)64 << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
__device__ __forceinline__ void mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}" ::"r"(d), "l"(a),
               "l"(b), "r"(idesc), "r"(acc)
               : "memory");
}
__global__ void __launch_bounds__(128, 1) chain(int n, int iters, int n_acc, int issuers, unsigned long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar[2];
  __shared__ uint32_t slot;
  for (int i = threadIdx.x; i < 48 * 1024 / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0x3c003c00u;
  const uint32_t b = smem_u32(&bar[threadIdx.x >> 5 & 1]);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[0])));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[1])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = slot;
  if ((threadIdx.x & 31) == 0 && (threadIdx.x >> 5) < issuers) {
    const uint32_t my = (threadIdx.x >> 5) * 256;
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | (8u << 24);
    const uint64_t ad = desc_sw128(smem_u32(smem)), bd = desc_sw128(smem_u32(smem) + 16384);
    const unsigned long long t0 = clock64();
    for (int i = 0; i < iters; ++i) mma(tmem + my + (uint32_t)((i % n_acc) * n), ad + 2 * (i & 3), bd + 2 * (i & 3), idesc, i >= n_acc);
    const unsigned long long t1 = clock64();
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(b) : "memory");
    uint32_t done = 0;
    while (!done)
      asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0,1,0,p;\n}" : "=r"(done) : "r"(b) : "memory");
    const unsigned long long t2 = clock64();
    out[2 * (threadIdx.x >> 5)] = t1 - t0;
    out[2 * (threadIdx.x >> 5) + 1] = t2 - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}
int main() {
  unsigned long long* d;
  cudaMalloc(&d, 32);
  cudaFuncSetAttribute(chain, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  const int ns[] = {16, 64, 128, 256};
  for (int n : ns)
    for (int issuers : {1, 2}) {
      const int iters = 512, n_acc = 1;
      chain<<<1, 128, 64 * 1024>>>(n, iters, n_acc, issuers, d);
      unsigned long long h[4];
      cudaError_t e = cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost);
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
      printf("M=128 N=%3d issuers=%d iters=%3d each: thread0 issue %6llu complete %6llu | thread32 issue %6llu complete %6llu -> %6.1f clk per MMA overall (floor %d)\n",
             n, issuers, iters, h[0], h[1], issuers > 1 ? h[2] : 0ull, issuers > 1 ? h[3] : 0ull,
             (double)(issuers > 1 && h[3] > h[1] ? h[3] : h[1]) / (iters * issuers), 128 * n / 256);
    }
  return 0;
}
