// Development micro-benchmark: per-SM turnaround between the exit of a CTA and the entry of the next
// kernel's CTA on the same SM, for back-to-back launches of a persistent-style kernel (one CTA per
// SM), as a function of dynamic shared memory, TMEM allocation, PDL and how long the kernel runs.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <algorithm>
#include <vector>

#include "ptx.cuh"
using namespace rnb;

#define CK(x)                                                                  \
  do {                                                                         \
    cudaError_t e_ = (x);                                                      \
    if (e_ != cudaSuccess) {                                                   \
      printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__);      \
      return 1;                                                                \
    }                                                                          \
  } while (0)

__device__ unsigned long long g_t[8][2][256];
__device__ unsigned int g_sm[8][256];

__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ unsigned long long g_dealloc[256];

__global__ void __launch_bounds__(320) k(int launch, int spin_us, int use_tmem, int touch_smem,
                                          float* sink, int store_bytes, int mma_count, int mma_n,
                                          int ld_count) {
  extern __shared__ uint8_t smem[];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(8) uint64_t mma_bar;
  if (threadIdx.x == 0) {
    ptx::mbar_init(ptx::smem_u32(&mma_bar), 1);
    ptx::fence_mbar_init();
  }
  if (threadIdx.x == 0) {
    g_t[launch & 7][0][blockIdx.x] = gtimer();
    unsigned int smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    g_sm[launch & 7][blockIdx.x] = smid;
  }
  ptx::grid_dep_launch_dependents();
  if (use_tmem && threadIdx.x < 32) {
    ptx::tmem_alloc(ptx::smem_u32(&tmem_slot), 512);
    ptx::tmem_relinquish();
  }
  if (touch_smem) {
    for (int i = threadIdx.x * 4; i < touch_smem; i += blockDim.x * 4)
      *reinterpret_cast<uint32_t*>(smem + i) = i;
  }
  ptx::grid_dep_wait();
  const unsigned long long t0 = gtimer();
  // optional stores, spread over the run
  size_t off = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 8;
  int stored = 0;
  while (gtimer() - t0 < (unsigned long long)spin_us * 1000ull) {
    if (stored < store_bytes) {
      float4 v = make_float4(1, 2, 3, 4);
      *reinterpret_cast<float4*>(sink + off) = v;
      *reinterpret_cast<float4*>(sink + off + 4) = v;
      off += (size_t)gridDim.x * blockDim.x * 8;
      stored += 32 * blockDim.x;
    }
  }
  __syncthreads();
  if (use_tmem && mma_count > 0 && threadIdx.x < 32) {
    // garbage operands from dynamic shared memory (needs >= 64 KB): only the timing matters
    ptx::tc_fence_after();
    const uint32_t base = tmem_slot;
    const uint32_t sbase = (ptx::smem_u32(smem) + 1023u) & ~1023u;
    const uint64_t adesc = ptx::make_kmajor_desc(sbase, 128);
    const uint64_t bdesc = ptx::make_kmajor_desc(sbase + 16384, 128);
    const uint32_t idesc = ptx::make_idesc_bf16_m128((uint32_t)mma_n);
    for (int i = 0; i < mma_count; ++i)
      ptx::umma_bf16_elect(base + (i & 1) * 256, adesc, bdesc, idesc, i > 1);
    ptx::umma_commit_elect(ptx::smem_u32(&mma_bar));
    ptx::mbar_wait(ptx::smem_u32(&mma_bar), 0);
    ptx::tc_fence_after();
  }
  __syncthreads();
  if (use_tmem && ld_count > 0 && threadIdx.x >= 64 && threadIdx.x < 192) {
    const uint32_t base = tmem_slot;
    const int w = (threadIdx.x >> 5) & 3;
    uint32_t v[16];
    uint32_t acc = 0;
    for (int i = 0; i < ld_count; ++i) {
      ptx::tmem_ld16(base + ((uint32_t)(w * 32) << 16) + (i & 15) * 16, v);
      ptx::tmem_ld_wait();
      acc += v[0];
    }
    if (acc == 0x12345u) sink[0] = 1.f;
    ptx::tc_fence_before();
  }
  __syncthreads();
  if (threadIdx.x == 0) g_t[launch & 7][1][blockIdx.x] = gtimer();
  if (use_tmem && threadIdx.x < 32) {
    ptx::tc_fence_after();
    uint32_t base = tmem_slot;
    const unsigned long long d0 = gtimer();
    ptx::tmem_dealloc(base, 512);
    if (threadIdx.x == 0) g_dealloc[blockIdx.x] = gtimer() - d0;
  }
}

int main() {
  int dev = 0, num_sms = 0;
  CK(cudaSetDevice(dev));
  CK(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev));
  CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
  float* sink;
  const size_t sink_bytes = 512ull << 20;
  CK(cudaMalloc(&sink, sink_bytes));
  struct Cfg {
    const char* name;
    int smem, spin_us, tmem, touch, pdl, store_kb, mma, n, ld;
  } cfgs[] = {
      {"tmem, 20us, 0 MMA", 200 * 1024, 20, 1, 0, 1, 0, 0, 256, 0},
      {"tmem, 20us, 100 MMA N=256", 200 * 1024, 20, 1, 0, 1, 0, 100, 256, 0},
      {"tmem, 20us, 800 MMA N=256", 200 * 1024, 20, 1, 0, 1, 0, 800, 256, 0},
      {"tmem, 20us, 800 MMA N=64", 200 * 1024, 20, 1, 0, 1, 0, 800, 64, 0},
      {"tmem, 20us, 0 MMA, 100 ld", 200 * 1024, 20, 1, 0, 1, 0, 0, 256, 100},
      {"tmem, 20us, 800 MMA N=256, 100 ld", 200 * 1024, 20, 1, 0, 1, 0, 800, 256, 100},
      {"tmem, 1us, 800 MMA N=256, 100 ld", 200 * 1024, 1, 1, 0, 1, 0, 800, 256, 100},
      {"smem 0, 20us", 0, 20, 0, 0, 1, 0},
      {"smem 200K, 20us", 200 * 1024, 20, 0, 0, 1, 0},
      {"smem 200K touched, 20us", 200 * 1024, 20, 0, 200 * 1024, 1, 0},
      {"smem 200K + tmem, 20us", 200 * 1024, 20, 1, 0, 1, 0},
      {"smem 200K + tmem, 40us", 200 * 1024, 40, 1, 0, 1, 0},
      {"smem 200K + tmem, 20us, no PDL", 200 * 1024, 20, 1, 0, 0, 0},
      {"smem 200K + tmem, 40us, 256 KB stores/CTA", 200 * 1024, 40, 1, 0, 1, 256},
      {"smem 200K + tmem, 40us, 768 KB stores/CTA", 200 * 1024, 40, 1, 0, 1, 768},
  };
  for (const Cfg& c : cfgs) {
    for (int l = 0; l < 8; ++l) {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(num_sms);
      cfg.blockDim = dim3(320);
      cfg.dynamicSmemBytes = c.smem;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[0].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = attr;
      cfg.numAttrs = c.pdl ? 1 : 0;
      CK(cudaLaunchKernelEx(&cfg, k, l, c.spin_us, c.tmem, c.touch, sink, c.store_kb * 1024, c.mma, c.n,
                             c.ld));
    }
    CK(cudaDeviceSynchronize());
    static unsigned long long t[8][2][256];
    static unsigned int sm[8][256];
    CK(cudaMemcpyFromSymbol(t, g_t, sizeof(t)));
    CK(cudaMemcpyFromSymbol(sm, g_sm, sizeof(sm)));
    std::vector<double> turn, period;
    for (int l = 4; l < 7; ++l) {
      for (int i = 0; i < num_sms; ++i)
        for (int j = 0; j < num_sms; ++j)
          if (sm[l + 1][j] == sm[l][i])
            turn.push_back((double)(long long)(t[l + 1][0][j] - t[l][1][i]) * 1e-3);
      unsigned long long a = ~0ull, b = ~0ull;
      for (int i = 0; i < num_sms; ++i) {
        a = std::min(a, t[l][0][i]);
        b = std::min(b, t[l + 1][0][i]);
      }
      period.push_back((double)(b - a) * 1e-3);
    }
    std::sort(turn.begin(), turn.end());
    static unsigned long long dl[256];
    CK(cudaMemcpyFromSymbol(dl, g_dealloc, sizeof(dl)));
    std::vector<double> dv;
    for (int i = 0; i < num_sms; ++i) dv.push_back(dl[i] * 1e-3);
    std::sort(dv.begin(), dv.end());
    printf("dealloc us med %.2f max %.2f | ", dv[dv.size() / 2], dv.back());
    printf("%-46s turnaround us: min %.2f med %.2f max %.2f   launch period %.1f us\n", c.name,
           turn.empty() ? 0 : turn[0], turn.empty() ? 0 : turn[turn.size() / 2],
           turn.empty() ? 0 : turn.back(), period[1]);
  }
  return 0;
}
