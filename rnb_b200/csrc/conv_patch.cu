// Patch convolution (see conv_patch.cuh): one activation patch per tile, filter taps as
// row-shifted UMMA descriptors, weights resident in shared memory or streamed through a ring,
// epilogue straight from TMEM to global memory.
//
//   warp 0      TMA producer: (resident mode) all weight tiles once; per work item one patch
//               (nkb boxes, double buffered) and (stream mode) num_taps x nkb weight tiles
//   warp 1      TMEM owner and MMA issuer.  The whole warp runs the issue loop converged and one
//               elected lane issues each tcgen05.mma: with a shuffle-broadcast warp index / TMEM
//               address and all-lane mbarrier waits, ptxas keeps every descriptor in uniform
//               registers (back-to-back UTCHMMA).  Running the role inside `if (lane == 0)`, or
//               waiting with one lane + __syncwarp, costs ~120-146 cycles of R2UR traffic per MMA
//               against 72 cycles of tensor time at N=144 (profiles/r01_mma_issue_notes.md).
//   warps 2-9   epilogue: tcgen05.ld -> +shift (+residual) -> ReLU -> bf16 -> 32-byte global stores
#include <stdio.h>

#include "conv_patch.cuh"
#include "ptx.cuh"

namespace rnb {

namespace {

struct PatchItem {
  int w0, h0, t0, b;  // position tile
  int n0, ncols;      // channel tile
};

// work items: channel tile fastest, then the position tiles in (w, h, t, b) order
__device__ __forceinline__ PatchItem decode_patch_item(const ConvPatchParams& p, int item) {
  PatchItem c;
  const int nt = item % p.n_tiles;
  int tile = item / p.n_tiles;
  const int iw = tile % p.tiles_w;
  tile /= p.tiles_w;
  const int ih = tile % p.tiles_h;
  tile /= p.tiles_h;
  const int it = tile % p.tiles_t;
  c.b = tile / p.tiles_t;
  c.w0 = iw * p.tw;
  c.h0 = ih * p.th;
  c.t0 = it * p.tt;
  c.n0 = nt * p.block_n;
  c.ncols = min(p.block_n, p.n - c.n0);
  return c;
}

}  // namespace

template <bool kRes, bool kRelu, bool kStream>
__global__ void __launch_bounds__(kConvThreads, 1)
conv_patch_kernel(const __grid_constant__ ConvPatchParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  // shuffle-broadcast so the compiler knows these are warp-uniform (role branches, TMEM address)
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  const uint32_t wtile_bytes = p.block_n * 128;         // one (tap, channel block) weight tile
  const uint32_t n_wtiles = p.num_taps * p.nkb;
  const uint32_t n_wslots = kStream ? (uint32_t)p.b_stages : n_wtiles;
  const uint32_t pblock_bytes = p.patch_rows * 128;     // one channel block of a patch
  const uint32_t patch_bytes = p.nkb * pblock_bytes;
  const uint32_t wbuf = smem_base;
  const uint32_t pbuf = wbuf + n_wslots * wtile_bytes;
  const uint32_t bars = pbuf + 2 * patch_bytes + 1024;  // +1024: rows past the last patch that
                                                        // only garbage accumulator rows read
  const uint32_t w_full = bars;
  const uint32_t p_full = bars + 8;    // [2]
  const uint32_t p_empty = bars + 24;  // [2]
  const uint32_t tmem_full_bar = bars + 40;   // [2]
  const uint32_t tmem_empty_bar = bars + 56;  // [2]
  const uint32_t tmem_ptr_smem = bars + 72;
  const uint32_t b_full = bars + 80;                          // [kPatchMaxBStages]
  const uint32_t b_empty = bars + 80 + 8 * kPatchMaxBStages;  // [kPatchMaxBStages]
  const uint32_t shift_tab = bars + 256;

  const int total_items = p.tiles_w * p.tiles_h * p.tiles_t * p.tiles_b * p.n_tiles;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&p.a_map);
    ptx::prefetch_tmap(&p.b_map);
    ptx::mbar_init(w_full, 1);
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(p_full + 8 * s, 1);
      ptx::mbar_init(p_empty + 8 * s, 1);
      ptx::mbar_init(tmem_full_bar + 8 * s, 1);
      ptx::mbar_init(tmem_empty_bar + 8 * s, kEpiThreads);
    }
    for (int s = 0; s < kPatchMaxBStages; ++s) {
      ptx::mbar_init(b_full + 8 * s, 1);
      ptx::mbar_init(b_empty + 8 * s, 1);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_ptr_smem, kTmemCols);
    ptx::tmem_relinquish();
  }
  if (!kStream && threadIdx.x >= 64) {
    // one channel tile for the whole kernel: its shift table is loaded once
    const int et = threadIdx.x - 64;
    if (et < p.n) {
      const float v = __ldg(p.shift + et);
      asm volatile("st.shared.f32 [%0], %1;\n" ::"r"(shift_tab + et * 4), "f"(v) : "memory");
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  ptx::grid_dep_launch_dependents();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];\n" : "=r"(tmem_base) : "r"(tmem_ptr_smem));
  tmem_base = __shfl_sync(0xffffffffu, tmem_base, 0);

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (warp-wide)
    if (!kStream) {
      // the packed weights are constants of the model: fetch them while the previous kernel in
      // the stream is still draining, and only then wait for its activations
      ptx::mbar_expect_tx_elect(w_full, n_wtiles * wtile_bytes);
      for (int tap = 0, i = 0; tap < p.num_taps; ++tap)
        for (int kb = 0; kb < p.nkb; ++kb, ++i)
          ptx::tma_load_2d_elect(wbuf + i * wtile_bytes, &p.b_map, w_full,
                                 tap * p.cin_pad + kb * 64, 0);
    }
    ptx::grid_dep_wait();
    int local = 0;
    uint32_t bs = 0, bphase = 0;
    for (int item = blockIdx.x; item < total_items; item += gridDim.x, ++local) {
      const PatchItem tc = decode_patch_item(p, item);
      const uint32_t pb = local & 1;
      const uint32_t ph = (local >> 1) & 1;
      ptx::mbar_wait(p_empty + 8 * pb, ph ^ 1);
      ptx::mbar_expect_tx_elect(p_full + 8 * pb, patch_bytes);
      for (int kb = 0; kb < p.nkb; ++kb)
        ptx::tma_load_5d_elect(pbuf + pb * patch_bytes + kb * pblock_bytes, &p.a_map,
                               p_full + 8 * pb, kb * 64, tc.w0 + p.aw, tc.h0 + p.ah, tc.t0 + p.at,
                               tc.b);
      if (kStream) {
        for (int tap = 0; tap < p.num_taps; ++tap)
          for (int kb = 0; kb < p.nkb; ++kb) {
            ptx::mbar_wait(b_empty + 8 * bs, bphase ^ 1);
            ptx::mbar_expect_tx_elect(b_full + 8 * bs, wtile_bytes);
            ptx::tma_load_2d_elect(wbuf + bs * wtile_bytes, &p.b_map, b_full + 8 * bs,
                                   tap * p.cin_pad + kb * 64, tc.n0);
            if (++bs == (uint32_t)p.b_stages) {
              bs = 0;
              bphase ^= 1;
            }
          }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (warp-wide)
    // (reads only shared memory and TMEM: ordered behind the producer's grid_dep_wait by the
    // patch-full barrier)
    if (!kStream) ptx::mbar_wait(w_full, 0);
    int local = 0;
    uint32_t bs = 0, bphase = 0;
    for (int item = blockIdx.x; item < total_items; item += gridDim.x, ++local) {
      const PatchItem tc = decode_patch_item(p, item);
      const uint32_t idesc = ptx::make_idesc_bf16_m128((uint32_t)tc.ncols);
      const uint32_t pb = local & 1;
      const uint32_t ph = (local >> 1) & 1;
      const uint32_t tmem_d = tmem_base + pb * kMaxBlockN;
      ptx::mbar_wait(tmem_empty_bar + 8 * pb, ph ^ 1);
      ptx::mbar_wait(p_full + 8 * pb, ph);
      ptx::tc_fence_after();
      uint32_t accumulate = 0;
      uint32_t wt = wbuf;
      for (int tap = 0; tap < p.num_taps; ++tap) {
        const uint32_t arow = pbuf + pb * patch_bytes + (uint32_t)p.tap_shift[tap] * 128u;
        for (int kb = 0; kb < p.nkb; ++kb) {
          if (kStream) {
            ptx::mbar_wait(b_full + 8 * bs, bphase);
            ptx::tc_fence_after();
            wt = wbuf + bs * wtile_bytes;
          }
          const uint64_t adesc = ptx::make_kmajor_desc(arow + kb * pblock_bytes, 128);
          const uint64_t bdesc = ptx::make_kmajor_desc(wt, 128);
          const int ksteps = min(64, p.cin_pad - kb * 64) >> 4;
          ptx::umma_bf16_elect(tmem_d, adesc, bdesc, idesc, accumulate);
          if (ksteps > 1) ptx::umma_bf16_elect(tmem_d, adesc + 2, bdesc + 2, idesc, 1);
          if (ksteps > 2) ptx::umma_bf16_elect(tmem_d, adesc + 4, bdesc + 4, idesc, 1);
          if (ksteps > 3) ptx::umma_bf16_elect(tmem_d, adesc + 6, bdesc + 6, idesc, 1);
          accumulate = 1;
          if (kStream) {
            ptx::umma_commit_elect(b_empty + 8 * bs);
            if (++bs == (uint32_t)p.b_stages) {
              bs = 0;
              bphase ^= 1;
            }
          } else {
            wt += wtile_bytes;
          }
        }
      }
      ptx::umma_commit_elect(p_empty + 8 * pb);
      ptx::umma_commit_elect(tmem_full_bar + 8 * pb);
    }
  } else {
    // ------------------------------------------------------------------ epilogue (256 threads)
    ptx::grid_dep_wait();  // residual reads and output stores must not pass the previous kernel
    const int ew = warp - 2;
    const int q = warp & 3;
    const int half = ew >> 2;
    const int et = threadIdx.x - 64;
    const int row = q * 32 + lane;
    const int rw = row & ((1 << p.lpw) - 1);
    const int rh = (row >> p.lpw) & ((1 << p.lph) - 1);
    const int rt = row >> (p.lpw + p.lph);
    const bool in_tile = (rw < p.tw) && (rh < p.th) && (rt < p.tt);
    int local = 0;
    int table_n0 = kStream ? -1 : 0;
    for (int item = blockIdx.x; item < total_items; item += gridDim.x, ++local) {
      const PatchItem tc = decode_patch_item(p, item);
      const uint32_t acc = local & 1;
      const uint32_t acc_phase = (local >> 1) & 1;
      const int w = tc.w0 + rw, h = tc.h0 + rh, t = tc.t0 + rt;
      const bool valid = in_tile && (w < p.Wo) && (h < p.Ho) && (t < p.To);
      const size_t off =
          ((((size_t)tc.b * p.To + t) * p.Ho + h) * p.Wo + w) * p.cout_pad + tc.n0;
      __nv_bfloat16* out_row = p.y + off;
      const __nv_bfloat16* res_row = kRes ? p.residual + off : nullptr;
      if (kStream && tc.n0 != table_n0) {
        // shift table of this channel tile; the barriers order it against the previous readers
        ptx::named_bar_sync(1, kEpiThreads);
        if (et < tc.ncols) {
          const float v = __ldg(p.shift + tc.n0 + et);
          asm volatile("st.shared.f32 [%0], %1;\n" ::"r"(shift_tab + et * 4), "f"(v) : "memory");
        }
        ptx::named_bar_sync(1, kEpiThreads);
        table_n0 = tc.n0;
      }
      ptx::mbar_wait(tmem_full_bar + 8 * acc, acc_phase);
      ptx::tc_fence_after();
      const uint32_t taddr_row = tmem_base + ((uint32_t)(q * 32) << 16) + acc * kMaxBlockN;
      const int nchunk = tc.ncols >> 4;
      for (int cchunk = half; cchunk < nchunk; cchunk += 2) {
        uint32_t v[16];
        const int col = cchunk * 16;
        ptx::tmem_ld16(taddr_row + col, v);
        uint64_t sh2[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];\n"
                       : "=l"(sh2[2 * j]), "=l"(sh2[2 * j + 1])
                       : "r"(shift_tab + (col + 4 * j) * 4));
        }
        uint32_t rwd[8];
        if (kRes) {
#pragma unroll
          for (int j = 0; j < 8; ++j) rwd[j] = 0u;
          if (valid) ptx::ld_global_nc_v8(res_row + col, rwd);
        }
        ptx::tmem_ld_wait();
        uint32_t o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          uint64_t x = ptx::add_f32x2(
              ptx::pack_f32x2(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1])), sh2[j]);
          if (kRes)
            x = ptx::add_f32x2(x, ptx::pack_f32x2(__uint_as_float(rwd[j] << 16),
                                                  __uint_as_float(rwd[j] & 0xffff0000u)));
          o[j] = ptx::cvt_bf16x2<kRelu>(x);
        }
        if (valid) ptx::st_global_v8(out_row + col, o);
      }
      ptx::tc_fence_before();
      ptx::mbar_arrive(tmem_empty_bar + 8 * acc);
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, kTmemCols);
  }
}

namespace {
template <bool kRes, bool kRelu, bool kStream>
cudaError_t launch_patch_variant(const ConvPatchParams& p, int grid, int smem, cudaStream_t stream) {
  static bool configured[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    e = cudaFuncSetAttribute(conv_patch_kernel<kRes, kRelu, kStream>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit);
    if (e != cudaSuccess) return e;
    configured[dev] = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kConvThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, conv_patch_kernel<kRes, kRelu, kStream>, p);
}

template <bool kStream>
cudaError_t launch_patch_mode(const ConvPatchParams& p, int grid, int smem, cudaStream_t stream) {
  const bool res = p.residual != nullptr, relu = p.relu != 0;
  if (res) return relu ? launch_patch_variant<true, true, kStream>(p, grid, smem, stream)
                       : launch_patch_variant<true, false, kStream>(p, grid, smem, stream);
  return relu ? launch_patch_variant<false, true, kStream>(p, grid, smem, stream)
              : launch_patch_variant<false, false, kStream>(p, grid, smem, stream);
}
}  // namespace

cudaError_t launch_conv_patch(const ConvPatchParams& p, int num_sms, cudaStream_t stream) {
  if (p.b_stages < 0 || p.b_stages > kPatchMaxBStages) return cudaErrorInvalidValue;
  const int smem = conv_patch_smem_bytes(p.nkb, p.patch_rows, p.num_taps, p.block_n, p.b_stages);
  const int total = p.tiles_w * p.tiles_h * p.tiles_t * p.tiles_b * p.n_tiles;
  const int grid = total < num_sms ? total : num_sms;
  return p.b_stages > 0 ? launch_patch_mode<true>(p, grid, smem, stream)
                        : launch_patch_mode<false>(p, grid, smem, stream);
}

}  // namespace rnb
