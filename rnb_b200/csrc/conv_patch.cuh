// "Patch" variant of the tcgen05 convolution for stride-1 1x3x3 and 3x1x1 layers.
//
// conv_tc_kernel re-reads the 128-row activation tile once per filter tap, so an SM has to pull
// taps x (A + B) bytes per tile through its ~68 B/clk inbound port (983 MB L2->SM per conv2 launch,
// profiles/r01_ncu_conv_tc_v1.md) -- that port, not the tensor core, bounds it.  Here a CTA loads
// ONE activation patch (tile + halo) per tile and every filter tap is the same patch read at a row
// offset, expressed only in the UMMA shared-memory descriptor.  (Measured on B200: the 128-byte
// swizzle is a function of the shared-memory ADDRESS bits, so a descriptor whose start address is
// shifted by whole 128-byte rows reads correctly with the base-offset field left at zero; filling
// that field with address bits 7-9 gives wrong results.  profiles/r01_test_conv_patch.log)
//
// Weights: RESIDENT mode keeps all packed weights in shared memory for the life of the CTA (conv2
// stage, stem temporal); STREAM mode (weights too large) tiles the output channels and streams the
// weight tiles of one channel tile through a ring, once per (position tile, channel tile).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "conv_tc.cuh"

namespace rnb {

constexpr int kPatchMaxKb = 4;      // 64-channel blocks per patch
constexpr int kPatchMaxBStages = 8;
constexpr int kPatchMaxBufs = 4;    // activation patches in flight

struct ConvPatchParams {
  alignas(64) CUtensorMap a_map;  // activation (C, W, H, T, B), box (64, pbox_w, pbox_h, pbox_t, 1)
  alignas(64) CUtensorMap b_map;  // packed weights (K_total, N), box (64, block_n)
  // TAIL block (resident, non-pair mode): when cin_pad = 64 * nkb + 16 or + 32, the trailing
  // channels get a block of their own with 32- / 64-byte rows and swizzle instead of a zero-filled
  // 128-byte one: a quarter / half of the shared memory and of the fill bandwidth of that block
  alignas(64) CUtensorMap a_map_tail;  // box (tail_ch, pbox_w, pbox_h, pbox_t, 1)
  alignas(64) CUtensorMap b_map_tail;  // box (tail_ch, block_n)
  // residual through shared memory (res_smem = 1; resident single-CTA mode, n = 64 or 128): the
  // producer warp fetches the tile's residual rows by TMA a tile ahead, so that no global-memory
  // latency sits in the epilogue between "accumulator full" and "accumulator free"
  alignas(64) CUtensorMap r_map;       // residual (C, W, H, T, B), box (64, tile extents, 1)
  const float* shift;
  const __nv_bfloat16* residual;
  __nv_bfloat16* y;     // output (B, To, Ho, Wo, cout_pad): written with plain 32-byte stores
  int nkb;              // full 64-channel blocks of the input (cin_pad / 64, rounded up if no tail)
  int tail_ch;          // 0, or the 16 / 32 trailing channels held in the tail block
  int cin_pad;
  int patch_rows;       // rows (positions) of one patch block = pbox_w * pbox_h * pbox_t
  int num_taps;
  int n;                // output channels (multiple of 16)
  int n_tiles, block_n; // channel tiling: resident mode has one tile of n channels
  int b_stages;         // 0: weights resident; > 0: ring depth of the streamed weight tiles
  int p_bufs;           // patch ring depth (2..kPatchMaxBufs)
#ifdef RNB_PATCH_COUNTERS
  int dbg_launch;       // which of the 4 timestamp sets this launch writes
#endif
  int res_smem;
  int pair;             // 1: CTA pairs (cta_group::2), each CTA holds half of every weight tile
  // 1: clips are taken last to first.  A kernel whose input was written first-to-last by the launch
  // before it finds the clips written LAST still in L2 (the 115 MB conv2 intermediate does not fit
  // as a whole), so consecutive launches alternate the direction; results do not depend on it.
  int reverse;
  // output tile: tw x th x tt positions, laid out in the accumulator with power-of-two pitches
  int tw, th, tt;
  int lpw, lph;         // log2 of the row pitches: row = (t << (lpw + lph)) + (h << lpw) + w
  int tiles_w, tiles_h, tiles_t, tiles_b;
  int aw, ah, at;       // patch origin relative to the tile origin (-1 where the filter extends)
  int Wo, Ho, To, B;
  int cout_pad;
  int relu;
  int16_t tap_shift[kMaxTaps];  // row offset of each tap inside the patch
};

// block_n here is the rows of a weight tile that ONE CTA holds (half the channel tile in pair mode)
inline int conv_patch_smem_bytes(int nkb, int patch_rows, int num_taps, int block_n, int b_stages,
                                 int p_bufs = 2, int tail_ch = 0, int res_blocks = 0) {
  const int wbytes = b_stages > 0 ? b_stages * block_n * 128
                                  : num_taps * block_n * (nkb * 128 + tail_ch * 2);
  // (res_blocks: 64-channel blocks of a residual tile, two tiles in flight)
  // 1024 alignment slack, 2048 control block (barrier areas, shift table, TMEM address)
  return 1024 + 2048 + wbytes + p_bufs * patch_rows * (nkb * 128 + tail_ch * 2) +
         1024 /*slack rows*/ + 2 * res_blocks * 16384;
}

cudaError_t launch_conv_patch(const ConvPatchParams& p, int num_sms, cudaStream_t stream);

// Co-scheduled SpatioTemporalConv (conv_patch.cu): `sp` = the 1 x k x k convolution, `tp` = the
// k x 1 x 1 convolution reading sp's output, both resident-weight single-CTA patch plans over the
// same batch.  n_sp of the grid's CTAs start on the spatial convolution, the others on the
// temporal one; temporal items are claimed `chunk` at a time from a grid-wide counter, so the
// spatial CTAs take their share of what is left when they are done.
// `flags`: device array of (flag_clips + 2) zero-initialised 64-bit counters, flag_clips >= batch,
// owned by the caller and shared by all such launches of a stream (each launch re-zeroes it).
bool conv_patch_dual_applicable(const ConvPatchParams& sp, const ConvPatchParams& tp);
cudaError_t launch_conv_patch_dual(const ConvPatchParams& sp, const ConvPatchParams& tp, int n_sp,
                                   int chunk, unsigned long long* flags, int flag_clips,
                                   int num_sms, cudaStream_t stream);

}  // namespace rnb
