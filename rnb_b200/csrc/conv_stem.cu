// Fused stem kernel (see conv_stem.cuh).
//
//   warp 0      TMA producer: both weight sets once, then one patch (16 KB) per tile, ring of 4
//   warp 1      MMA issuer (warp-wide, one elected lane issues): spatial accumulator S of tile i+2,
//               then the temporal accumulator Y of tile i -- so the tensor core works on the next
//               tiles' 4x4 taps while the intermediate of this tile is being converted
//   warps 2-5   "mid" epilogue: S (TMEM) -> ReLU -> bf16 -> shared-memory operand of the temporal
//               convolution (128-byte-swizzled block of channels 0..63, 64-byte-swizzled block of
//               channels 64..95), proxy fence, hand over to the MMA warp
//   warps 6-9   output epilogue: Y (TMEM) -> bf16 -> 32-byte global stores
//   (the BatchNorm shift and the bias are weight rows of the two GEMMs, see conv_stem.cuh)
#include "conv_stem.cuh"

#include <atomic>
#include <stdio.h>

#include "conv_tc.cuh"
#include "ptx.cuh"

namespace rnb {

namespace {

constexpr uint32_t kWspTile = kStemMid * 32;        // one 4x4 tap: 96 rows of 16 channels
constexpr uint32_t kWtpTile = kStemOut * 128;       // one temporal tap, channels 0..63
constexpr uint32_t kWtpTail = kStemOut * 64;        // one temporal tap, channels 64..95
constexpr uint32_t kPatchBytes = 4 * 128 * 32;      // 4 s2d rows x 8 frames x 16 columns x 32 B
constexpr uint32_t kMidRows = (kStemT + 2) * 16;    // 160: one zero frame on either side
constexpr uint32_t kMidMain = kMidRows * 128;
constexpr uint32_t kMidTail = kMidRows * 64;
constexpr int kPatchBufs = 4;
constexpr int kSBufs = 3;  // spatial accumulators in flight (3 x 128 TMEM columns)

struct StemItem {
  int w0, ho, b;
};
__device__ __forceinline__ StemItem stem_decode(int item) {
  StemItem s;
  s.w0 = (item % kStemTilesW) * kStemTw;
  item /= kStemTilesW;
  s.ho = item % kStemHW;
  s.b = item / kStemHW;
  return s;
}

}  // namespace

__global__ void __launch_bounds__(kConvThreads, 1)
conv_stem_fused_kernel(const __grid_constant__ StemFusedParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  // control block
  const uint32_t w_full = base;
  const uint32_t p_full = base + 8;                 // [kPatchBufs]
  const uint32_t p_empty = p_full + 8 * kPatchBufs;  // [kPatchBufs]
  const uint32_t s_full = p_empty + 8 * kPatchBufs;  // [kSBufs]
  const uint32_t s_empty = s_full + 8 * kSBufs;
  const uint32_t m_full = s_empty + 8 * kSBufs;
  const uint32_t m_empty = m_full + 16;
  const uint32_t y_full = m_empty + 16;
  const uint32_t y_empty = y_full + 16;
  const uint32_t tmem_ptr_smem = base + 256;
  // operands
  const uint32_t wsp = base + 1024;
  const uint32_t wtp = wsp + 16 * kWspTile;
  const uint32_t wtp_tail = wtp + 3 * kWtpTile;
  const uint32_t patch = wtp_tail + 3 * kWtpTail;
  // (taps read up to 3 rows past a patch; the intermediate's 128-byte-swizzled block needs 1024)
  const uint32_t mid = patch + kPatchBufs * kPatchBytes + 1024;
  static_assert((16 * kWspTile) % 1024 == 0 && (3 * kWtpTile) % 1024 == 0 &&
                    (3 * kWtpTail) % 1024 == 0 && kPatchBytes % 1024 == 0 &&
                    kMidMain % 1024 == 0 && (kMidMain + kMidTail) % 1024 == 0,
                "operand alignment");

  const int total_items = p.B * kStemHW * kStemTilesW;
  const int first_item = blockIdx.x;
  const int item_step = gridDim.x;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&p.in_map);
    ptx::prefetch_tmap(&p.wsp_map);
    ptx::prefetch_tmap(&p.wtp_map);
    ptx::prefetch_tmap(&p.wtp_tail_map);
    ptx::mbar_init(w_full, 1);
    for (int s = 0; s < kPatchBufs; ++s) {
      ptx::mbar_init(p_full + 8 * s, 1);
      ptx::mbar_init(p_empty + 8 * s, 1);
    }
    for (int s = 0; s < kSBufs; ++s) {
      ptx::mbar_init(s_full + 8 * s, 1);
      ptx::mbar_init(s_empty + 8 * s, 4);  // one arrival per mid-epilogue warp
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(m_full + 8 * s, 4);
      ptx::mbar_init(m_empty + 8 * s, 1);
      ptx::mbar_init(y_full + 8 * s, 1);
      ptx::mbar_init(y_empty + 8 * s, 4);  // one arrival per output-epilogue warp
    }
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_ptr_smem, kTmemCols);
    ptx::tmem_relinquish();
  }
  // the zero frames of both intermediate buffers (never written again)
  {
    const uint32_t zero = 0u;
    // per buffer: main block rows 0..15 and 144..159 (2 x 2048 B), tail block the same (2 x 1024 B)
    for (int i = threadIdx.x; i < 2 * (2 * 512 + 2 * 256); i += blockDim.x) {
      const int buf = i / (2 * 512 + 2 * 256);
      int k = i % (2 * 512 + 2 * 256);
      uint32_t addr = mid + buf * (kMidMain + kMidTail);
      if (k < 1024) {
        addr += (k < 512 ? 0u : 144u * 128u - 2048u) + (uint32_t)k * 4u;
      } else {
        k -= 1024;
        addr += kMidMain + (k < 256 ? 0u : 144u * 64u - 1024u) + (uint32_t)k * 4u;
      }
      asm volatile("st.shared.b32 [%0], %1;\n" ::"r"(addr), "r"(zero) : "memory");
    }
  }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  ptx::grid_dep_launch_dependents();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];\n" : "=r"(tmem_base) : "r"(tmem_ptr_smem));
  tmem_base = __shfl_sync(0xffffffffu, tmem_base, 0);
  const uint32_t tmem_s = tmem_base;                 // kSBufs x 128 columns
  const uint32_t tmem_y = tmem_base + kSBufs * 128;  // 2 x 64 columns
  static_assert(kSBufs * 128 + 2 * 64 <= kTmemCols, "TMEM budget");

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (warp-wide)
    ptx::mbar_expect_tx_elect(w_full, 16 * kWspTile + 3 * kWtpTile + 3 * kWtpTail);
    for (int tap = 0; tap < 16; ++tap)
      ptx::tma_load_2d_elect(wsp + tap * kWspTile, &p.wsp_map, w_full, tap * 16, 0);
    for (int dt = 0; dt < 3; ++dt) {
      ptx::tma_load_2d_elect(wtp + dt * kWtpTile, &p.wtp_map, w_full, dt * kStemMid, 0);
      ptx::tma_load_2d_elect(wtp_tail + dt * kWtpTail, &p.wtp_tail_map, w_full, dt * kStemMid + 64, 0);
    }
    ptx::grid_dep_wait();
    uint32_t pb = 0, ph = 0;
    for (int item = first_item; item < total_items; item += item_step) {
      const StemItem it = stem_decode(item);
      ptx::mbar_wait(p_empty + 8 * pb, ph ^ 1);
      ptx::mbar_expect_tx_elect(p_full + 8 * pb, kPatchBytes);
      // dims (c, w, t, h, b): columns w0 - 2 .. w0 + 13, all frames, s2d rows ho - 2 .. ho + 1;
      // coordinates outside the image are zero-filled = the convolution's padding
      ptx::tma_load_5d_elect(patch + pb * kPatchBytes, &p.in_map, p_full + 8 * pb, 0, it.w0 - 2, 0,
                             it.ho - 2, it.b);
      if (++pb == kPatchBufs) {
        pb = 0;
        ph ^= 1;
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (warp-wide)
    const uint32_t idesc_s = ptx::make_idesc_bf16_m128(kStemMid);
    const uint32_t idesc_y = ptx::make_idesc_bf16_m128(kStemOut);
    int n_items = 0;
    for (int item = first_item; item < total_items; item += item_step) ++n_items;
    uint32_t pb = 0, ph = 0;
    auto spatial = [&](int i) {
      const uint32_t a = (uint32_t)i % kSBufs, aph = ((uint32_t)i / kSBufs) & 1;
      ptx::mbar_wait(s_empty + 8 * a, aph ^ 1);
      ptx::mbar_wait(p_full + 8 * pb, ph);
      ptx::tc_fence_after();
      const uint32_t pa = patch + pb * kPatchBytes;
      const uint32_t d = tmem_s + a * 128;
#pragma unroll 1
      for (int ah = 0; ah < 4; ++ah)
#pragma unroll
        for (int aw = 0; aw < 4; ++aw) {
          const uint64_t adesc = ptx::make_kmajor_desc(pa + (uint32_t)(ah * 128 + aw) * 32u, 32);
          const uint64_t bdesc = ptx::make_kmajor_desc(wsp + (uint32_t)(ah * 4 + aw) * kWspTile, 32);
          ptx::umma_bf16_elect(d, adesc, bdesc, idesc_s, (ah | aw) ? 1u : 0u);
        }
      ptx::umma_commit_elect(p_empty + 8 * pb);
      ptx::umma_commit_elect(s_full + 8 * a);
      if (++pb == kPatchBufs) {
        pb = 0;
        ph ^= 1;
      }
    };
    if (n_items > 0) {
      ptx::mbar_wait(w_full, 0);
      ptx::tc_fence_after();
      spatial(0);
      if (n_items > 1) spatial(1);
    }
    // order on the tensor pipe: S(0) S(1) | S(2) T(0) | S(3) T(1) | ...: the conversion of tile i
    // has two spatial and one temporal pass of other tiles to hide behind.  (Doing all the barrier
    // waits of an iteration first and then queueing both passes straight-line was measured slower:
    // 48.6 against 42.3 us.)
    for (int i = 0; i < n_items; ++i) {
      if (i + 2 < n_items) spatial(i + 2);
      const uint32_t a = i & 1, aph = (i >> 1) & 1;
      ptx::mbar_wait(m_full + 8 * a, aph);
      ptx::mbar_wait(y_empty + 8 * a, aph ^ 1);
      ptx::tc_fence_after();
      const uint32_t m0 = mid + a * (kMidMain + kMidTail);
      const uint32_t d = tmem_y + a * 64;
#pragma unroll 1
      for (int dt = 0; dt < 3; ++dt) {
        // tap dt = the rows one frame (16 rows) further on
        const uint64_t ad = ptx::make_kmajor_desc(m0 + (uint32_t)dt * 16u * 128u, 128);
        const uint64_t bd = ptx::make_kmajor_desc(wtp + (uint32_t)dt * kWtpTile, 128);
        ptx::umma_bf16_elect(d, ad, bd, idesc_y, dt ? 1u : 0u);
        ptx::umma_bf16_elect(d, ad + 2, bd + 2, idesc_y, 1);
        ptx::umma_bf16_elect(d, ad + 4, bd + 4, idesc_y, 1);
        ptx::umma_bf16_elect(d, ad + 6, bd + 6, idesc_y, 1);
        const uint64_t at = ptx::make_kmajor_desc(m0 + kMidMain + (uint32_t)dt * 16u * 64u, 64);
        const uint64_t bt = ptx::make_kmajor_desc(wtp_tail + (uint32_t)dt * kWtpTail, 64);
        ptx::umma_bf16_elect(d, at, bt, idesc_y, 1);
        ptx::umma_bf16_elect(d, at + 2, bt + 2, idesc_y, 1);
      }
      ptx::umma_commit_elect(m_empty + 8 * a);
      ptx::umma_commit_elect(y_full + 8 * a);
    }
  } else if (warp < 6) {
    // ------------------------------------------------------------------ mid epilogue (128 threads)
    const int q = warp & 3;
    const int row = q * 32 + lane;        // accumulator row = (t, column)
    const uint32_t mr = (uint32_t)row + 16;  // row of the intermediate operand (frame t + 1)
    int i = 0;
    for (int item = first_item; item < total_items; item += item_step, ++i) {
      const uint32_t sa = (uint32_t)i % kSBufs, sph = ((uint32_t)i / kSBufs) & 1;
      const uint32_t a = i & 1, aph = (i >> 1) & 1;
      ptx::mbar_wait(s_full + 8 * sa, sph);
      ptx::mbar_wait(m_empty + 8 * a, aph ^ 1);
      ptx::tc_fence_after();
      const uint32_t taddr = tmem_s + ((uint32_t)(q * 32) << 16) + sa * 128;
      const uint32_t m0 = mid + a * (kMidMain + kMidTail);
#pragma unroll
      for (int c = 0; c < kStemMid / 16; c += 2) {
        uint32_t va[16], vb[16];
        ptx::tmem_ld16(taddr + c * 16, va);
        ptx::tmem_ld16(taddr + (c + 1) * 16, vb);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const uint32_t(&v)[16] = h ? vb : va;
          const int cc = c + h;  // 16-channel chunk = two 16-byte units
          // (the folded BatchNorm shift is part of the GEMM: a constant-one input channel times a
          // weight row of shifts, see conv_stem.cuh) -> ReLU -> bf16
          uint32_t o[8];
#pragma unroll
          for (int j = 0; j < 8; ++j)
            o[j] = ptx::cvt_bf16x2<true>(
                ptx::pack_f32x2(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1])));
          // channel 95 (padding) carries the constant one that multiplies the temporal bias row
          if (cc == kStemMid / 16 - 1) o[7] = (o[7] & 0x0000ffffu) | 0x3f800000u;
          uint32_t a0, a1;
          if (cc < 4) {  // channels 0..63: 128-byte rows, unit u at u ^ (row & 7)
            const uint32_t rb = m0 + mr * 128u, x7 = mr & 7u;
            a0 = rb + (((uint32_t)(2 * cc) ^ x7) << 4);
            a1 = rb + (((uint32_t)(2 * cc + 1) ^ x7) << 4);
          } else {       // channels 64..95: 64-byte rows, unit u at u ^ ((row >> 1) & 3)
            const uint32_t rb = m0 + kMidMain + mr * 64u, x3 = (mr >> 1) & 3u;
            a0 = rb + (((uint32_t)(2 * (cc - 4)) ^ x3) << 4);
            a1 = rb + (((uint32_t)(2 * (cc - 4) + 1) ^ x3) << 4);
          }
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};\n" ::"r"(a0), "r"(o[0]), "r"(o[1]),
                       "r"(o[2]), "r"(o[3])
                       : "memory");
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};\n" ::"r"(a1), "r"(o[4]), "r"(o[5]),
                       "r"(o[6]), "r"(o[7])
                       : "memory");
        }
      }
      // the tensor core reads the operand through the async proxy
      ptx::fence_proxy_async_smem();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        ptx::mbar_arrive(m_full + 8 * a);
        ptx::mbar_arrive(s_empty + 8 * sa);
      }
    }
  } else {
    // ------------------------------------------------------------------ output epilogue (128 threads)
    ptx::grid_dep_wait();  // the output slot may still be read by the previous forward's kernels
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int t = row >> 4, j = row & 15;
    int i = 0;
    for (int item = first_item; item < total_items; item += item_step, ++i) {
      const StemItem it = stem_decode(item);
      const uint32_t a = i & 1, aph = (i >> 1) & 1;
      const bool valid = j < kStemTw && it.w0 + j < kStemHW;
      __nv_bfloat16* out =
          p.y + ((((size_t)it.b * kStemT + t) * kStemHW + it.ho) * kStemHW + it.w0 + j) * kStemOut;
      ptx::mbar_wait(y_full + 8 * a, aph);
      ptx::tc_fence_after();
      const uint32_t taddr = tmem_y + ((uint32_t)(q * 32) << 16) + a * 64;
#pragma unroll
      for (int c = 0; c < kStemOut / 16; c += 2) {
        uint32_t va[16], vb[16];
        ptx::tmem_ld16(taddr + c * 16, va);
        ptx::tmem_ld16(taddr + (c + 1) * 16, vb);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const uint32_t(&v)[16] = h ? vb : va;
          const int cc = c + h;
          uint32_t o[8];
#pragma unroll
          for (int k = 0; k < 8; ++k)
            o[k] = ptx::cvt_bf16x2<false>(
                ptx::pack_f32x2(__uint_as_float(v[2 * k]), __uint_as_float(v[2 * k + 1])));
          if (valid) ptx::st_global_v8(out + cc * 16, o);
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(y_empty + 8 * a);
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, kTmemCols);
  }
}

cudaError_t launch_conv_stem_fused(const StemFusedParams& p, int num_sms, cudaStream_t stream) {
  static std::atomic<bool> configured[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    e = cudaFuncSetAttribute(conv_stem_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             kStemFusedSmem);
    if (e != cudaSuccess) return e;
    configured[dev] = true;
  }
  const int items = p.B * kStemHW * kStemTilesW;
  if (items <= 0) return cudaSuccess;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(items < num_sms ? items : num_sms);
  cfg.blockDim = dim3(kConvThreads);
  cfg.dynamicSmemBytes = kStemFusedSmem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[3];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = add_l2_output_window(attr, 1, p.y,
                                      (size_t)p.B * kStemT * kStemHW * kStemHW * kStemOut * 2);
  return cudaLaunchKernelEx(&cfg, conv_stem_fused_kernel, p);
}

}  // namespace rnb
