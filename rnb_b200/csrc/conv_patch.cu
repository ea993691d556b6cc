// Patch convolution (see conv_patch.cuh): resident weights, one activation patch per tile, filter
// taps as row-shifted UMMA descriptors, epilogue straight from TMEM to global memory.
//
//   warp 0      loads all weight tiles once, then one patch (nkb boxes) per tile, double buffered
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

struct PatchTile {
  int w0, h0, t0, b;
};

__device__ __forceinline__ PatchTile decode_patch_tile(const ConvPatchParams& p, int tile) {
  PatchTile c;
  const int iw = tile % p.tiles_w;
  tile /= p.tiles_w;
  const int ih = tile % p.tiles_h;
  tile /= p.tiles_h;
  const int it = tile % p.tiles_t;
  c.b = tile / p.tiles_t;
  c.w0 = iw * p.tw;
  c.h0 = ih * p.th;
  c.t0 = it * p.tt;
  return c;
}

}  // namespace

template <bool kRes, bool kRelu>
__global__ void __launch_bounds__(kConvThreads, 1)
conv_patch_kernel(const __grid_constant__ ConvPatchParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  // shuffle-broadcast so the compiler knows these are warp-uniform (role branches, TMEM address)
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  const uint32_t wtile_bytes = p.n * 128;               // one (tap, channel block) weight tile
  const uint32_t n_wtiles = p.num_taps * p.nkb;
  const uint32_t pblock_bytes = p.patch_rows * 128;     // one channel block of a patch
  const uint32_t patch_bytes = p.nkb * pblock_bytes;
  const uint32_t wbuf = smem_base;
  const uint32_t pbuf = wbuf + n_wtiles * wtile_bytes;
  const uint32_t bars = pbuf + 2 * patch_bytes + 1024;  // +1024: rows past the last patch that
                                                        // only garbage accumulator rows read
  const uint32_t w_full = bars;
  const uint32_t p_full = bars + 8;    // [2]
  const uint32_t p_empty = bars + 24;  // [2]
  const uint32_t tmem_full_bar = bars + 40;   // [2]
  const uint32_t tmem_empty_bar = bars + 56;  // [2]
  const uint32_t tmem_ptr_smem = bars + 72;
  const uint32_t shift_tab = bars + 256;

  const int total_tiles = p.tiles_w * p.tiles_h * p.tiles_t * p.tiles_b;

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
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_ptr_smem, kTmemCols);
    ptx::tmem_relinquish();
  }
  if (threadIdx.x >= 64) {
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
    // the packed weights are constants of the model: fetch them while the previous kernel in the
    // stream is still draining, and only then wait for its activations
    ptx::mbar_expect_tx_elect(w_full, n_wtiles * wtile_bytes);
    for (int tap = 0, i = 0; tap < p.num_taps; ++tap)
      for (int kb = 0; kb < p.nkb; ++kb, ++i)
        ptx::tma_load_2d_elect(wbuf + i * wtile_bytes, &p.b_map, w_full, tap * p.cin_pad + kb * 64, 0);
    ptx::grid_dep_wait();
    int local = 0;
    long long wait_a = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++local) {
      const PatchTile tc = decode_patch_tile(p, tile);
      const uint32_t pb = local & 1;
      const uint32_t ph = (local >> 1) & 1;
      const long long tw0 = clock64();
      ptx::mbar_wait(p_empty + 8 * pb, ph ^ 1);
      wait_a += clock64() - tw0;
      ptx::mbar_expect_tx_elect(p_full + 8 * pb, patch_bytes);
      for (int kb = 0; kb < p.nkb; ++kb)
        ptx::tma_load_5d_elect(pbuf + pb * patch_bytes + kb * pblock_bytes, &p.a_map,
                               p_full + 8 * pb, kb * 64, tc.w0 + p.aw, tc.h0 + p.ah, tc.t0 + p.at,
                               tc.b);
    }
    if (p.dbg && lane == 0) p.dbg[blockIdx.x * 8 + 0] = wait_a;
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (warp-wide)
    // (reads only shared memory and TMEM: ordered behind the producer's grid_dep_wait by the
    // patch-full barrier)
    const uint32_t idesc = ptx::make_idesc_bf16_m128((uint32_t)p.n);
    const long long tm0 = clock64();
    ptx::mbar_wait(w_full, 0);
    long long wait_w = clock64() - tm0, wait_e = 0, wait_p = 0;
    int local = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++local) {
      const uint32_t pb = local & 1;
      const uint32_t ph = (local >> 1) & 1;
      const uint32_t tmem_d = tmem_base + pb * kMaxBlockN;
      const long long t0 = clock64();
      ptx::mbar_wait(tmem_empty_bar + 8 * pb, ph ^ 1);
      const long long t1 = clock64();
      ptx::mbar_wait(p_full + 8 * pb, ph);
      const long long t2 = clock64();
      wait_e += t1 - t0;
      wait_p += t2 - t1;
      ptx::tc_fence_after();
      uint32_t accumulate = 0;
      uint32_t wt = wbuf;
      for (int tap = 0; tap < p.num_taps; ++tap) {
        const uint32_t arow = pbuf + pb * patch_bytes + (uint32_t)p.tap_shift[tap] * 128u;
        for (int kb = 0; kb < p.nkb; ++kb, wt += wtile_bytes) {
          const uint64_t adesc = ptx::make_kmajor_desc(arow + kb * pblock_bytes, 128);
          const uint64_t bdesc = ptx::make_kmajor_desc(wt, 128);
          const int ksteps = min(64, p.cin_pad - kb * 64) >> 4;
          ptx::umma_bf16_elect(tmem_d, adesc, bdesc, idesc, accumulate);
          if (ksteps > 1) ptx::umma_bf16_elect(tmem_d, adesc + 2, bdesc + 2, idesc, 1);
          if (ksteps > 2) ptx::umma_bf16_elect(tmem_d, adesc + 4, bdesc + 4, idesc, 1);
          if (ksteps > 3) ptx::umma_bf16_elect(tmem_d, adesc + 6, bdesc + 6, idesc, 1);
          accumulate = 1;
        }
      }
      ptx::umma_commit_elect(p_empty + 8 * pb);
      ptx::umma_commit_elect(tmem_full_bar + 8 * pb);
    }
    if (p.dbg && lane == 0) {
      p.dbg[blockIdx.x * 8 + 1] = wait_w;
      p.dbg[blockIdx.x * 8 + 2] = wait_e;
      p.dbg[blockIdx.x * 8 + 3] = wait_p;
      p.dbg[blockIdx.x * 8 + 4] = clock64() - tm0;
      p.dbg[blockIdx.x * 8 + 5] = local;
    }
  } else {
    // ------------------------------------------------------------------ epilogue (256 threads)
    ptx::grid_dep_wait();  // residual reads and output stores must not pass the previous kernel
    const int ew = warp - 2;
    const int q = warp & 3;
    const int half = ew >> 2;
    const int row = q * 32 + lane;
    const int rw = row & ((1 << p.lpw) - 1);
    const int rh = (row >> p.lpw) & ((1 << p.lph) - 1);
    const int rt = row >> (p.lpw + p.lph);
    const bool in_tile = (rw < p.tw) && (rh < p.th) && (rt < p.tt);
    const int nchunk = p.n >> 4;
    long long wait_f = 0;
    const long long te0 = clock64();
    int local = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++local) {
      const PatchTile tc = decode_patch_tile(p, tile);
      const uint32_t acc = local & 1;
      const uint32_t acc_phase = (local >> 1) & 1;
      const int w = tc.w0 + rw, h = tc.h0 + rh, t = tc.t0 + rt;
      const bool valid = in_tile && (w < p.Wo) && (h < p.Ho) && (t < p.To);
      const size_t off = ((((size_t)tc.b * p.To + t) * p.Ho + h) * p.Wo + w) * p.cout_pad;
      __nv_bfloat16* out_row = p.y + off;
      const __nv_bfloat16* res_row = kRes ? p.residual + off : nullptr;
      const long long tf0 = clock64();
      ptx::mbar_wait(tmem_full_bar + 8 * acc, acc_phase);
      wait_f += clock64() - tf0;
      ptx::tc_fence_after();
      const uint32_t taddr_row = tmem_base + ((uint32_t)(q * 32) << 16) + acc * kMaxBlockN;
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
    if (p.dbg && threadIdx.x == 64) {
      p.dbg[blockIdx.x * 8 + 6] = wait_f;
      p.dbg[blockIdx.x * 8 + 7] = clock64() - te0;
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
template <bool kRes, bool kRelu>
cudaError_t launch_patch_variant(const ConvPatchParams& p, int grid, int smem, cudaStream_t stream) {
  static bool configured[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    e = cudaFuncSetAttribute(conv_patch_kernel<kRes, kRelu>,
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
  return cudaLaunchKernelEx(&cfg, conv_patch_kernel<kRes, kRelu>, p);
}
}  // namespace

cudaError_t launch_conv_patch(const ConvPatchParams& p, int num_sms, cudaStream_t stream) {
  const int smem = conv_patch_smem_bytes(p.nkb, p.patch_rows, p.num_taps, p.n);
  const int total_tiles = p.tiles_w * p.tiles_h * p.tiles_t * p.tiles_b;
  const int grid = total_tiles < num_sms ? total_tiles : num_sms;
  const bool res = p.residual != nullptr, relu = p.relu != 0;
  if (res) return relu ? launch_patch_variant<true, true>(p, grid, smem, stream)
                       : launch_patch_variant<true, false>(p, grid, smem, stream);
  return relu ? launch_patch_variant<false, true>(p, grid, smem, stream)
              : launch_patch_variant<false, false>(p, grid, smem, stream);
}

}  // namespace rnb
