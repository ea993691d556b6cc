// Micro-benchmark: cycles per tcgen05.mma (cta_group::1, M=128, K=16, bf16, operands in smem) for
// several N, with aligned and row-shifted A descriptors.  Dev tool; not part of the library.
#include <stdio.h>
#include <stdlib.h>
#include "ptx.cuh"
using namespace rnb;

__global__ void __launch_bounds__(128, 1) mma_rate(int n, int iters, int shift_rows, int distinct,
                                                   long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  __shared__ uint64_t bar;
  __shared__ uint32_t tptr;
  // zero operands (finite)
  for (int i = threadIdx.x; i < 200 * 1024 / 16; i += blockDim.x)
    asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" ::"r"(base + i * 16), "r"(0) : "memory");
  if (threadIdx.x == 0) {
    ptx::mbar_init(ptx::smem_u32(&bar), 1);
    ptx::fence_mbar_init();
  }
  if (threadIdx.x < 32) {
    ptx::tmem_alloc(ptx::smem_u32(&tptr), 512);
    ptx::tmem_relinquish();
  }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  if (threadIdx.x == 0) {
    const uint32_t idesc = ptx::make_idesc_bf16_m128(n);
    const uint32_t a0 = base, b0 = base + 96 * 1024;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      // 36 MMAs like one conv2 tile: 9 "taps" x 4 k-steps
      for (int tap = 0; tap < 9; ++tap) {
        const uint32_t a = a0 + (distinct ? (tap % 3) * 128 * shift_rows + (tap / 3) * 2048 : 0);
        const uint32_t b = b0 + (distinct ? tap * 2048 : 0);
        const uint64_t ad = ptx::make_kmajor_desc(a, 128), bd = ptx::make_kmajor_desc(b, 128);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          ptx::umma_bf16(tptr, ad + 2 * k, bd + 2 * k, idesc, (tap | k) ? 1u : 0u);
      }
    }
    ptx::umma_commit(ptx::smem_u32(&bar));
    long long t1 = clock64();
    ptx::mbar_wait(ptx::smem_u32(&bar), 0);
    long long t2 = clock64();
    if (blockIdx.x == 0) {
      out[0] = t1 - t0;
      out[1] = t2 - t0;
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) ptx::tmem_dealloc(tptr, 512);
}

int main() {
  long long* d;
  cudaMalloc(&d, 16);
  cudaFuncSetAttribute(mma_rate, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  const int iters = 200;
  for (int grid : {1, 148}) {
    for (int n : {64, 128, 144, 192, 256}) {
      for (int mode = 0; mode < 3; ++mode) {  // 0: same operands, 1: distinct aligned, 2: distinct shifted by 1 row
        mma_rate<<<grid, 128, 220 * 1024>>>(n, iters, mode == 2 ? 1 : 0, mode > 0, d);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
        long long h[2];
        cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        const double per = (double)h[1] / (iters * 36.0);
        printf("grid %3d N=%3d mode %d: issue %.1f cyc/mma, complete %.1f cyc/mma  -> %.0f MAC/cyc/SM (ideal N/2 = %d cyc)\n",
               grid, n, mode, (double)h[0] / (iters * 36.0), per, 128.0 * n * 16 / per, n / 2);
      }
    }
  }
  return 0;
}
