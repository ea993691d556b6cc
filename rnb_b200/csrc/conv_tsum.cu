// Temporal convolution with the taps stacked along N and summed in the epilogue (conv_tsum.cuh).
//
//   warp 0      TMA producer: the weights once (per 64-channel block the three taps' [64 x 64]
//               tiles one under the other = a [192 x 64] K-major operand), then one activation
//               tile (128 rows x cin_pad channels, no halo) per work item, ring of p_bufs
//   warp 1      MMA issuer (warp-wide, one elected lane issues): cin_pad / 16 MMAs of N = 192
//   warps 2-9   epilogue: three 16-column chunks (one per tap) from TMEM, the outer two shifted by
//               one lane (= one frame), + shift (+ residual) -> ReLU -> bf16 -> 32-byte stores
//   The residual rows of a tile come in by TMA (warp 0) into a ring of their own, laid out like
//   the accumulator rows, a few tiles ahead of the epilogue.
#include "conv_tsum.cuh"

#include <atomic>
#include <stdio.h>

#include "epilogue.cuh"
#include "ptx.cuh"

namespace rnb {

namespace {

constexpr uint32_t kAccStride = 256;  // TMEM columns between the two accumulators (192 used)

template <bool kRes, bool kRelu>
__global__ void __launch_bounds__(kConvThreads, 1)
conv_tsum_kernel(const __grid_constant__ ConvTsumParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  // control block
  const uint32_t w_full = base;
  const uint32_t p_full = base + 8;                       // [p_bufs <= 8]
  const uint32_t p_empty = p_full + 64;                   // [p_bufs]
  const uint32_t tmem_full_bar = p_empty + 64;            // [2]
  const uint32_t tmem_empty_bar = tmem_full_bar + 16;     // [2]
  const uint32_t r_full = tmem_empty_bar + 16;            // [r_bufs <= 4]
  const uint32_t r_empty = r_full + 32;                   // [r_bufs]
  const uint32_t tmem_ptr_smem = base + 240;
  const uint32_t shift_tab = base + 256;                  // 64 floats
  // operands
  const uint32_t tail_row = (uint32_t)p.tail_ch * 2u;     // bytes per row of the tail blocks
  const uint32_t w_main = base + 1024;                    // [nkb][192 rows][128 B]
  const uint32_t w_tail = w_main + (uint32_t)p.nkb * (kTsumN * 128u);
  const uint32_t patches = w_tail + kTsumN * tail_row;
  const uint32_t patch_bytes = (uint32_t)p.nkb * 16384u + 128u * tail_row;
  const uint32_t res_bufs = patches + (uint32_t)p.p_bufs * patch_bytes;  // [r_bufs][128 rows][128 B]

  const int total_items = p.B * p.tiles_hw;
  const int first_item = blockIdx.x;
  const int item_step = gridDim.x;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&p.a_map);
    ptx::prefetch_tmap(&p.b_map);
    if (p.tail_ch) {
      ptx::prefetch_tmap(&p.a_map_tail);
      ptx::prefetch_tmap(&p.b_map_tail);
    }
    if (kRes) ptx::prefetch_tmap(&p.r_map);
    ptx::mbar_init(w_full, 1);
    for (int s = 0; s < p.p_bufs; ++s) {
      ptx::mbar_init(p_full + 8 * s, 1);
      ptx::mbar_init(p_empty + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(tmem_full_bar + 8 * s, 1);
      ptx::mbar_init(tmem_empty_bar + 8 * s, 8);  // one arrival per epilogue warp
    }
    for (int s = 0; s < p.r_bufs; ++s) {
      ptx::mbar_init(r_full + 8 * s, 1);
      ptx::mbar_init(r_empty + 8 * s, 8);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_ptr_smem, kTmemCols);
    ptx::tmem_relinquish();
  }
  if (threadIdx.x >= 64 && threadIdx.x < 64 + kTsumCout) {
    // (constant data of the layer, not produced by the previous launch: read before the wait)
    const float v = p.shift[threadIdx.x - 64];
    asm volatile("st.shared.f32 [%0], %1;\n" ::"r"(shift_tab + (threadIdx.x - 64) * 4u), "f"(v)
                 : "memory");
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  ptx::grid_dep_launch_dependents();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];\n" : "=r"(tmem_base) : "r"(tmem_ptr_smem));
  tmem_base = __shfl_sync(0xffffffffu, tmem_base, 0);

  auto clip_of = [&](int item) {
    const int b = item / p.tiles_hw;
    return p.reverse ? p.B - 1 - b : b;
  };

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (warp-wide)
    ptx::mbar_expect_tx_elect(w_full, kTsumN * ((uint32_t)p.nkb * 128u + tail_row));
    for (int kb = 0; kb < p.nkb; ++kb)
      for (int tap = 0; tap < 3; ++tap)
        ptx::tma_load_2d_elect(w_main + (uint32_t)(kb * 3 + tap) * 8192u, &p.b_map, w_full,
                               tap * p.cin_pad + kb * 64, 0);
    if (p.tail_ch)
      for (int tap = 0; tap < 3; ++tap)
        ptx::tma_load_2d_elect(w_tail + (uint32_t)tap * kTsumCout * tail_row, &p.b_map_tail, w_full,
                               tap * p.cin_pad + p.nkb * 64, 0);
    ptx::grid_dep_wait();
    uint32_t pb = 0, ph = 0, rs = 0, rph = 0;
    for (int item = first_item; item < total_items; item += item_step) {
      int hw0 = (item % p.tiles_hw) * kTsumPos, b = clip_of(item);
#ifdef RNB_PATCH_COUNTERS
      if (p.dbg & 1) hw0 = (item % 8) * kTsumPos, b = 0;
#endif
      const uint32_t dst = patches + pb * patch_bytes;
      ptx::mbar_wait(p_empty + 8 * pb, ph ^ 1);
      ptx::mbar_expect_tx_elect(p_full + 8 * pb, patch_bytes);
      // dims (c, t, hw, b, 1): positions past HW are zero-filled (their rows are never stored)
      for (int kb = 0; kb < p.nkb; ++kb)
        ptx::tma_load_5d_elect(dst + (uint32_t)kb * 16384u, &p.a_map, p_full + 8 * pb, kb * 64, 0,
                               hw0, b, 0);
      if (p.tail_ch)
        ptx::tma_load_5d_elect(dst + (uint32_t)p.nkb * 16384u, &p.a_map_tail, p_full + 8 * pb,
                               p.nkb * 64, 0, hw0, b, 0);
      if (++pb == (uint32_t)p.p_bufs) {
        pb = 0;
        ph ^= 1;
      }
      if (kRes) {
        // the tile's residual rows, laid out like the accumulator rows, for the epilogue
        ptx::mbar_wait(r_empty + 8 * rs, rph ^ 1);
        ptx::mbar_expect_tx_elect(r_full + 8 * rs, 16384u);
        ptx::tma_load_5d_elect(res_bufs + rs * 16384u, &p.r_map, r_full + 8 * rs, 0, 0,
                               (item % p.tiles_hw) * kTsumPos, clip_of(item), 0);
        if (++rs == (uint32_t)p.r_bufs) {
          rs = 0;
          rph ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (warp-wide)
    const uint32_t idesc = ptx::make_idesc_bf16_m128(kTsumN);
    uint32_t pb = 0, ph = 0;
    int local = 0;
    for (int item = first_item; item < total_items; item += item_step, ++local) {
      const uint32_t acc = local & 1, acc_phase = (local >> 1) & 1;
      const uint32_t d = tmem_base + acc * kAccStride;
      if (local == 0) {
        ptx::mbar_wait(w_full, 0);
        ptx::tc_fence_after();
      }
      ptx::mbar_wait(tmem_empty_bar + 8 * acc, acc_phase ^ 1);
      ptx::mbar_wait(p_full + 8 * pb, ph);
      ptx::tc_fence_after();
      const uint32_t pa = patches + pb * patch_bytes;
      uint32_t accumulate = 0;
#ifdef RNB_PATCH_COUNTERS
      if (!(p.dbg & 8))
#endif
      for (int kb = 0; kb < p.nkb; ++kb) {
        const uint64_t ad = ptx::make_kmajor_desc(pa + (uint32_t)kb * 16384u, 128);
        const uint64_t bd = ptx::make_kmajor_desc(w_main + (uint32_t)kb * (kTsumN * 128u), 128);
        ptx::umma_bf16_elect(d, ad, bd, idesc, accumulate);
        ptx::umma_bf16_elect(d, ad + 2, bd + 2, idesc, 1);
        ptx::umma_bf16_elect(d, ad + 4, bd + 4, idesc, 1);
        ptx::umma_bf16_elect(d, ad + 6, bd + 6, idesc, 1);
        accumulate = 1;
      }
      if (tail_row) {
        const uint64_t ad = ptx::make_kmajor_desc(pa + (uint32_t)p.nkb * 16384u, tail_row);
        const uint64_t bd = ptx::make_kmajor_desc(w_tail, tail_row);
        ptx::umma_bf16_elect(d, ad, bd, idesc, accumulate);
        if (tail_row > 32) ptx::umma_bf16_elect(d, ad + 2, bd + 2, idesc, 1);
      }
      ptx::umma_commit_elect(p_empty + 8 * pb);
      ptx::umma_commit_elect(tmem_full_bar + 8 * acc);
      if (++pb == (uint32_t)p.p_bufs) {
        pb = 0;
        ph ^= 1;
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (256 threads)
    // Software-pipelined over (tile, chunk) units: while a chunk is summed, converted and stored
    // the TMEM loads of the next one are in flight, and the accumulator is handed back to the MMA
    // warp as soon as both chunks of a tile are in registers.  (Sixteen epilogue warps with one
    // chunk each were measured much slower, 52 against 31 us: the shuffles of the tap sum go
    // through the same data pipe as the operand reads of the MMAs and the TMA fills.)
    ptx::grid_dep_wait();  // output stores must not pass the previous kernel
    const int ew = warp - 2;
    const int q = warp & 3;     // TMEM lane quarter this warp may read
    const int half = ew >> 2;   // chunks half and half + 2 of the four 16-channel chunks
    const int row = q * 32 + lane;
    const int t = row & 7, pos = row >> 3;
    const bool has_prev = t > 0, has_next = t < kTsumT - 1;
    const uint32_t lane_taddr = tmem_base + ((uint32_t)(q * 32) << 16);
    auto issue = [&](uint32_t acc, int c, uint32_t(&v0)[16], uint32_t(&v1)[16], uint32_t(&v2)[16]) {
      const uint32_t ta = lane_taddr + acc * kAccStride + (uint32_t)c * 16u;
      ptx::tmem_ld16(ta, v0);
      ptx::tmem_ld16(ta + kTsumCout, v1);
      ptx::tmem_ld16(ta + 2 * kTsumCout, v2);
    };
    auto finish = [&](const uint32_t(&v0)[16], const uint32_t(&v1)[16], const uint32_t(&v2)[16],
                      int c, long long off, uint32_t res_tile) {
      uint64_t sh[8];
      epi::load_shift16(shift_tab, c * 16, sh);
      uint32_t r[8];
      if (kRes) epi::load_res16_smem(res_tile, row, c, r);
      uint32_t s[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        // tap 0 reads frame t - 1 (the lane below), tap 2 frame t + 1 (the lane above)
        const float up = __shfl_up_sync(0xffffffffu, __uint_as_float(v0[j]), 1);
        const float dn = __shfl_down_sync(0xffffffffu, __uint_as_float(v2[j]), 1);
        float x = __uint_as_float(v1[j]);
        x += has_prev ? up : 0.f;
        x += has_next ? dn : 0.f;
        s[j] = __float_as_uint(x);
      }
      uint32_t o[8];
      epi::finish16<kRes, kRelu>(s, sh, r, o);
#ifdef RNB_PATCH_COUNTERS
      if (p.dbg & 2) {
        if (o[0] == 0x12345678u && o[7] == 0x9abcdef0u) ptx::st_global_v8(p.y + off + c * 16, o);
        return;
      }
#endif
      if (off >= 0) ptx::st_global_v8(p.y + off + c * 16, o);
    };
    int n_items = 0;
    for (int item = first_item; item < total_items; item += item_step) ++n_items;
    uint32_t a0[16], a1[16], a2[16], b0[16], b1[16], b2[16];
    if (n_items > 0) {
      ptx::mbar_wait(tmem_full_bar, 0);
      ptx::tc_fence_after();
      issue(0, half, a0, a1, a2);
    }
    uint32_t rs = 0, rph = 0;
    for (int local = 0; local < n_items; ++local) {
      const int item = first_item + local * item_step;
      const uint32_t acc = local & 1;
      const int hw = (item % p.tiles_hw) * kTsumPos + pos;
      const long long off =
          hw < p.HW ? (((long long)clip_of(item) * kTsumT + t) * p.HW + hw) * kTsumCout : -1;
      const uint32_t res_tile = res_bufs + rs * 16384u;
#ifdef RNB_PATCH_COUNTERS
      const bool math = !(p.dbg & 4);
#else
      const bool math = true;
#endif
      // ---- chunk `half`; the loads of chunk `half + 2` go out behind it
      ptx::tmem_ld_wait();
      issue(acc, half + 2, b0, b1, b2);
      if (kRes) ptx::mbar_wait(r_full + 8 * rs, rph);
      if (math) finish(a0, a1, a2, half, off, res_tile);
      // ---- chunk `half + 2`: with it in registers the accumulator is free for the tile after next
      ptx::tmem_ld_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(tmem_empty_bar + 8 * acc);
      bool issued = true;
      if (local + 1 < n_items) {
        const uint32_t nacc = (local + 1) & 1, nph = ((local + 1) >> 1) & 1;
        // the next tile's first chunk now if its accumulator is already complete (the usual
        // case), otherwise after this chunk's arithmetic
        issued = __all_sync(0xffffffffu, ptx::mbar_try_wait(tmem_full_bar + 8 * nacc, nph));
        if (issued) {
          ptx::tc_fence_after();
          issue(nacc, half, a0, a1, a2);
        }
      }
      if (math) finish(b0, b1, b2, half + 2, off, res_tile);
      if (kRes) {
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(r_empty + 8 * rs);
        if (++rs == (uint32_t)p.r_bufs) {
          rs = 0;
          rph ^= 1;
        }
      }
      if (!issued) {
        const uint32_t nacc = (local + 1) & 1, nph = ((local + 1) >> 1) & 1;
        ptx::mbar_wait(tmem_full_bar + 8 * nacc, nph);
        ptx::tc_fence_after();
        issue(nacc, half, a0, a1, a2);
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, kTmemCols);
  }
}

template <bool kRes, bool kRelu>
cudaError_t launch_variant(const ConvTsumParams& p, int num_sms, int smem, cudaStream_t stream) {
  static std::atomic<bool> configured[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    e = cudaFuncSetAttribute(conv_tsum_kernel<kRes, kRelu>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit);
    if (e != cudaSuccess) return e;
    configured[dev] = true;
  }
  const int items = p.B * p.tiles_hw;
  if (items <= 0) return cudaSuccess;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(items < num_sms ? items : num_sms);
  cfg.blockDim = dim3(kConvThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[3];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = add_l2_output_window(attr, 1, p.y, (size_t)p.B * kTsumT * p.HW * kTsumCout * 2);
  return cudaLaunchKernelEx(&cfg, conv_tsum_kernel<kRes, kRelu>, p);
}

}  // namespace

cudaError_t launch_conv_tsum(const ConvTsumParams& p, int num_sms, cudaStream_t stream) {
  if (p.nkb < 0 || p.nkb > kTsumMaxKb || p.p_bufs < 2 || p.p_bufs > 8 || p.r_bufs < 0 ||
      p.r_bufs > 4 || (p.residual != nullptr) != (p.r_bufs > 0) ||
      (p.tail_ch != 0 && p.tail_ch != 16 && p.tail_ch != 32) || p.nkb * 64 + p.tail_ch < 16)
    return cudaErrorInvalidValue;
  const int smem = conv_tsum_smem_bytes(p.nkb, p.tail_ch, p.p_bufs, p.r_bufs);
  if (smem > kSmemLimit) return cudaErrorInvalidValue;
  const bool res = p.residual != nullptr, relu = p.relu != 0;
  if (res) return relu ? launch_variant<true, true>(p, num_sms, smem, stream)
                       : launch_variant<true, false>(p, num_sms, smem, stream);
  return relu ? launch_variant<false, true>(p, num_sms, smem, stream)
              : launch_variant<false, false>(p, num_sms, smem, stream);
}

}  // namespace rnb
