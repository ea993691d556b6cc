// Epilogue shared by the two tcgen05 convolution kernels: accumulator (TMEM) -> + folded BatchNorm
// shift -> (+ residual) -> (ReLU) -> bf16 -> global, one output position (= TMEM lane) per thread.
//
// Memory-level parallelism is what matters here (profiles/r01_patch_counters.md): a thread handles
// two 16-channel chunks per iteration, the residual rows of the first iteration are requested
// BEFORE the wait for the accumulator and those of the next iteration before the arithmetic of the
// current one, so that no DRAM latency sits between "accumulator full" and "accumulator free".
#pragma once
#include <cuda_bf16.h>
#include <stdint.h>

#include "ptx.cuh"

#ifdef RNB_PATCH_COUNTERS
#define EPI_MARK(k)                                       \
  {                                                       \
    const long long epi_now = clock64();                  \
    epi_cnt[k] += (unsigned long long)(epi_now - epi_t);  \
    epi_t = epi_now;                                      \
  }
#else
#define EPI_MARK(k)
#endif

namespace rnb {

namespace epi {
__device__ __forceinline__ void load_shift16(uint32_t shift_tab, int col, uint64_t (&sh)[8]) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];\n"
                 : "=l"(sh[2 * j]), "=l"(sh[2 * j + 1])
                 : "r"(shift_tab + (col + 4 * j) * 4));
  }
}
template <bool kRes>
__device__ __forceinline__ void load_res16(const __nv_bfloat16* res_row, int col, bool pred,
                                           uint32_t (&r)[8]) {
  if (kRes) {
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = 0u;
    if (pred) ptx::ld_global_nc_v8(res_row + col, r);
  }
}
// Residual chunk c (16 channels) of accumulator row `row` from a TMA-written shared-memory tile:
// blocks of 64 channels, [128 rows][128 bytes] each, 128-byte swizzle (16-byte unit u of row r sits
// at unit u ^ (r & 7)).
__device__ __forceinline__ void load_res16_smem(uint32_t res_tile, int row, int c, uint32_t (&r)[8]) {
  const uint32_t base = res_tile + (uint32_t)(c >> 2) * 16384u + (uint32_t)row * 128u;
  const uint32_t u0 = (uint32_t)(c & 3) * 2u, x = (uint32_t)row & 7u;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(base + ((u0 ^ x) << 4)));
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];\n"
               : "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(base + (((u0 + 1u) ^ x) << 4)));
}
template <bool kRes, bool kRelu>
__device__ __forceinline__ void finish16(const uint32_t (&v)[16], const uint64_t (&sh)[8],
                                         const uint32_t (&r)[8], uint32_t (&o)[8]) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    uint64_t x = ptx::add_f32x2(
        ptx::pack_f32x2(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1])), sh[j]);
    if (kRes)
      x = ptx::add_f32x2(x, ptx::pack_f32x2(__uint_as_float(r[j] << 16),
                                            __uint_as_float(r[j] & 0xffff0000u)));
    o[j] = ptx::cvt_bf16x2<kRelu>(x);
  }
}
}  // namespace epi

// L2 prefetch of the residual row segment [0, ncols) of one output position.  It has no register
// destination: a global LOAD kept in flight across the conversion of the previous tile makes that
// tile's tcgen05.ld / ld.shared results wait for it (the hardware scoreboards count in order), which
// costs more than the DRAM latency it hides (profiles/r01_patch_counters.md).  The caller issues
// this a whole tile ahead; the real loads of the tile then hit L2.
__device__ __forceinline__ void prefetch_residual_l2(const __nv_bfloat16* res_row, int ncols,
                                                     bool valid) {
  if (valid)
    for (int c = 0; c < ncols; c += 64)  // 128-byte lines
      asm volatile("prefetch.global.L2 [%0];\n" ::"l"(res_row + c));
}

// One accumulator tile.  `taddr_row` = TMEM address of this warp's lane quarter in the accumulator
// buffer, `nchunk` = 16-column chunks in the channel tile, `half` (0/1, warp-uniform) = which of the
// alternating chunks this warp converts, `valid` = this thread's output position exists.  Waits on
// `full_bar` (phase `full_phase`) itself, after the residual rows of the first two chunks have been
// requested; later chunks are requested one iteration ahead.
template <bool kRes, bool kRelu>
__device__ __forceinline__ void epilogue_tile(uint32_t full_bar, uint32_t full_phase,
                                              uint32_t taddr_row, uint32_t shift_tab, int nchunk,
                                              int half, bool valid, __nv_bfloat16* out_row,
                                              const __nv_bfloat16* res_row
#ifdef RNB_PATCH_COUNTERS
                                              ,
                                              unsigned long long (&epi_cnt)[6]
#endif
                                              ,
                                              uint32_t res_tile = 0, int row = 0,
                                              uint32_t res_bar = 0, uint32_t res_phase = 0) {
#ifdef RNB_PATCH_COUNTERS
  long long epi_t = clock64();
#endif
  uint32_t ra[8], rb[8];
  const bool res_smem = kRes && res_tile != 0;
  if (!res_smem) {
    epi::load_res16<kRes>(res_row, half * 16, valid && half < nchunk, ra);
    epi::load_res16<kRes>(res_row, (half + 2) * 16, valid && half + 2 < nchunk, rb);
  }
  EPI_MARK(0);
  ptx::mbar_wait(full_bar, full_phase);
  ptx::tc_fence_after();
  if (res_smem) {
    ptx::mbar_wait(res_bar, res_phase);
    if (half < nchunk) epi::load_res16_smem(res_tile, row, half, ra);
    if (half + 2 < nchunk) epi::load_res16_smem(res_tile, row, half + 2, rb);
  }
  EPI_MARK(1);
  for (int c0 = half; c0 < nchunk; c0 += 4) {
    const int c1 = c0 + 2;
    const bool two = c1 < nchunk;  // warp-uniform: tcgen05.ld is a warp-collective
    uint32_t va[16], vb[16];
    ptx::tmem_ld16(taddr_row + c0 * 16, va);
    if (two) ptx::tmem_ld16(taddr_row + c1 * 16, vb);
    EPI_MARK(5);
    uint64_t sa[8], sb[8];
    epi::load_shift16(shift_tab, c0 * 16, sa);
    if (two) epi::load_shift16(shift_tab, c1 * 16, sb);
    // the residual rows of the next iteration, while this one converts
    uint32_t na[8], nb[8];
    if (res_smem) {
      if (c0 + 4 < nchunk) epi::load_res16_smem(res_tile, row, c0 + 4, na);
      if (c1 + 4 < nchunk) epi::load_res16_smem(res_tile, row, c1 + 4, nb);
    } else {
      epi::load_res16<kRes>(res_row, (c0 + 4) * 16, valid && c0 + 4 < nchunk, na);
      epi::load_res16<kRes>(res_row, (c1 + 4) * 16, valid && c1 + 4 < nchunk, nb);
    }
    EPI_MARK(2);
    ptx::tmem_ld_wait();
    EPI_MARK(3);
    uint32_t o[8];
    epi::finish16<kRes, kRelu>(va, sa, ra, o);
    if (valid) ptx::st_global_v8(out_row + c0 * 16, o);
    if (two) {
      epi::finish16<kRes, kRelu>(vb, sb, rb, o);
      if (valid) ptx::st_global_v8(out_row + c1 * 16, o);
    }
    if (kRes) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        ra[j] = na[j];
        rb[j] = nb[j];
      }
    }
    EPI_MARK(4);
  }
}

}  // namespace rnb
