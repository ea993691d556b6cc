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
  const float* shift;
  const __nv_bfloat16* residual;
  __nv_bfloat16* y;     // output (B, To, Ho, Wo, cout_pad): written with plain 32-byte stores
  int nkb;              // 64-channel blocks of the input (cin_pad / 64 rounded up)
  int cin_pad;
  int patch_rows;       // rows (positions) of one patch block = pbox_w * pbox_h * pbox_t
  int num_taps;
  int n;                // output channels (multiple of 16)
  int n_tiles, block_n; // channel tiling: resident mode has one tile of n channels
  int b_stages;         // 0: weights resident; > 0: ring depth of the streamed weight tiles
  int p_bufs;           // patch ring depth (2..kPatchMaxBufs)
  int prefetch;         // L2-prefetch the patch this many items ahead (0 = off)
#ifdef RNB_PATCH_COUNTERS
  int dbg_launch;       // which of the 4 timestamp sets this launch writes
#endif
  int pair;             // 1: CTA pairs (cta_group::2), each CTA holds half of every weight tile
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
                                 int p_bufs = 2) {
  const int wtiles = b_stages > 0 ? b_stages : num_taps * nkb;
  return 1024 + wtiles * block_n * 128 + p_bufs * nkb * patch_rows * 128 + 1024 /*slack rows*/ +
         256 + kMaxBlockN * 4;
}

cudaError_t launch_conv_patch(const ConvPatchParams& p, int num_sms, cudaStream_t stream);

}  // namespace rnb
