// Micro-benchmark: aggregate L2 -> shared-memory bandwidth of TMA tile loads (2-D boxes of
// rows x 128 B from an L2-resident bf16 matrix), all SMs, ring of `stages` buffers, no compute.
// Dev tool; not part of the library.
#include <stdio.h>
#include <stdlib.h>
#include <string>
#include "plan.cuh"
#include "ptx.cuh"
using namespace rnb;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

__global__ void __launch_bounds__(64, 1)
tma_rate(const __grid_constant__ CUtensorMap map, int rows_per_box, int stages, int iters, int total_rows,
         long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t box_bytes = rows_per_box * 128;
  const uint32_t bars = base + stages * box_bytes;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) ptx::mbar_init(bars + 8 * s, 1);
    ptx::fence_mbar_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const long long t0 = clock64();
    uint32_t row = (blockIdx.x * 9973u * rows_per_box) % (uint32_t)total_rows;
    const uint32_t step = rows_per_box * 37, limit = (uint32_t)total_rows - rows_per_box;
    int s = 0;
    uint32_t phase = 0;
    // prologue: fill the ring
    for (int i = 0; i < stages; ++i) {
      ptx::mbar_expect_tx(bars + 8 * i, box_bytes);
      ptx::tma_load_2d(base + i * box_bytes, &map, bars + 8 * i, 0, (int)row);
      row += step;
      if (row > limit) row -= limit;
    }
    for (int i = 0; i < iters; ++i) {
      ptx::mbar_wait(bars + 8 * s, phase);
      ptx::mbar_expect_tx(bars + 8 * s, box_bytes);
      ptx::tma_load_2d(base + s * box_bytes, &map, bars + 8 * s, 0, (int)row);
      row += step;
      if (row > limit) row -= limit;
      if (++s == stages) {
        s = 0;
        phase ^= 1;
      }
    }
    out[blockIdx.x] = clock64() - t0;
  }
}

int main() {
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, 0);
  const int total_rows = 256 * 1024;  // 32 MB of 128-byte rows: L2 resident
  void* buf;
  cudaMalloc(&buf, (size_t)total_rows * 128);
  cudaMemset(buf, 0, (size_t)total_rows * 128);
  long long* d;
  cudaMalloc(&d, 148 * 8);
  void* fnp = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fnp, cudaEnableDefault, &q);
  EncodeTiledFn enc = (EncodeTiledFn)fnp;
  cudaFuncSetAttribute(tma_rate, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int rows : {64, 128, 256}) {
    CUtensorMap map;
    cuuint64_t gdim[2] = {64, (cuuint64_t)total_rows}, gstr[1] = {128};
    cuuint32_t box[2] = {64, (cuuint32_t)rows}, estr[2] = {1, 1};
    CUresult r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, buf, gdim, gstr, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
    for (int stages : {2, 4, 8, 12}) {
      if (stages * rows * 128 > 200 * 1024) continue;
      const int iters = 4000;
      for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(e0);
        tma_rate<<<prop.multiProcessorCount, 64, 210 * 1024>>>(map, rows, stages, iters, total_rows, d);
        cudaEventRecord(e1);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
      }
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      const double bytes = (double)prop.multiProcessorCount * iters * rows * 128;
      printf("box %3d rows (%5.1f KB) x %2d stages: %.2f TB/s aggregate, %.1f B/clk/SM at %d MHz nominal\n",
             rows, rows * 128 / 1024.0, stages, bytes / (ms * 1e-3) / 1e12,
             bytes / prop.multiProcessorCount / (ms * 1e-3) / (prop.clockRate * 1e3), prop.clockRate / 1000);
    }
  }
  return 0;
}
