// Micro-benchmark: cycles per tcgen05.mma.cta_group::2 (M=256, K=16, bf16) issued by the leader of a
// CTA pair, operands (zeros) in both CTAs' shared memory.  Dev tool; not part of the library.
#include <stdio.h>
#include <stdlib.h>
#include "ptx.cuh"
using namespace rnb;

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
mma_pair_rate(int n, int iters, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  __shared__ uint64_t bar;
  __shared__ uint32_t tptr;
  const uint32_t rank = ptx::cluster_ctarank();
  for (int i = threadIdx.x; i < 200 * 1024 / 16; i += blockDim.x)
    asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" ::"r"(base + i * 16), "r"(0) : "memory");
  if (threadIdx.x == 0) {
    ptx::mbar_init(ptx::smem_u32(&bar), 1);
    ptx::fence_mbar_init();
  }
  if (threadIdx.x < 32) {
    ptx::tmem_alloc_pair(ptx::smem_u32(&tptr), 512);
    ptx::tmem_relinquish_pair();
  }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync_all();
  ptx::tc_fence_after();
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  uint32_t tb = __shfl_sync(0xffffffffu, tptr, 0);
  if (warp == 1 && rank == 0) {
    const uint32_t idesc = ptx::make_idesc_bf16_m256(n);
    const uint32_t a0 = base, b0 = base + 96 * 1024;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      for (int tap = 0; tap < 9; ++tap) {
        const uint32_t a = a0 + (tap % 3) * 128 + (tap / 3) * 2048;
        const uint32_t b = b0 + tap * 2048;
        const uint64_t ad = ptx::make_kmajor_desc(a, 128), bd = ptx::make_kmajor_desc(b, 128);
        ptx::umma_bf16_pair_elect(tb, ad, bd, idesc, (it | tap) ? 1u : 0u);
        ptx::umma_bf16_pair_elect(tb, ad + 2, bd + 2, idesc, 1u);
        ptx::umma_bf16_pair_elect(tb, ad + 4, bd + 4, idesc, 1u);
        ptx::umma_bf16_pair_elect(tb, ad + 6, bd + 6, idesc, 1u);
      }
    }
    ptx::umma_commit_pair_elect(ptx::smem_u32(&bar), 1);
    ptx::mbar_wait(ptx::smem_u32(&bar), 0);
    long long t2 = clock64();
    if (blockIdx.x == 0 && (threadIdx.x & 31) == 0) out[0] = t2 - t0;
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync_all();
  if (threadIdx.x < 32) ptx::tmem_dealloc_pair(tb, 512);
}

int main() {
  long long* d;
  cudaMalloc(&d, 16);
  cudaFuncSetAttribute(mma_pair_rate, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  const int iters = 200;
  for (int grid : {2, 148}) {
    for (int n : {64, 128, 144, 192, 256}) {
      mma_pair_rate<<<grid, 128, 220 * 1024>>>(n, iters, d);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
      long long h[2];
      cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
      const double per = (double)h[0] / (iters * 36.0);
      printf("grid %3d pair-MMA M=256 N=%3d: %.1f cyc/mma -> %.0f MAC/cyc/SM (1-CTA floor N/2 = %d cyc)\n",
             grid, n, per, 256.0 * n * 16 / per / 2, n / 2);
    }
  }
  return 0;
}
