// Temporal 3x1x1 convolution (stride 1, padding 1) with the filter taps as OUTPUT columns.
//
// conv_patch.cu runs a 3x1x1 layer as 3 taps x K/16 tcgen05.mma of N = cout = 64 columns, each
// re-reading its 128 activation rows from shared memory: at N = 64 an MMA is bound by the operand
// reads ((128 + 64) rows x 32 B against a 128 B/clk port, 48 cycles) and the activation rows cross
// that port three times per tile (measured on the conv2 layers: 68 cycles per MMA, 27 MMAs per
// tile, profiles/r02_conv_tsum.md).  Here the three taps are stacked along N instead:
//
//     P[row, tap * 64 + co] = sum_ci x[row, ci] * w[tap, ci, co]          (one GEMM, N = 192)
//     y[t, pos, co]         = P[(t-1, pos), 0 * 64 + co] + P[(t, pos), 64 + co] + P[(t+1, pos), 128 + co]
//
// so a tile is K/16 MMAs of N = 192 (96 cycles each, tensor-bound) and every activation row is
// read once.  The tap sum happens in the epilogue: a tile holds ALL frames of 16 positions with
// the frame index fastest (row = pos * 8 + t, a TMA box whose dimensions are ordered (c, t, hw)),
// so the rows (t-1, pos) and (t+1, pos) are the neighbouring TMEM lanes = the neighbouring threads
// of the same warp, one __shfl_up / __shfl_down away; the frames before the first and after the
// last are the convolution's zero padding and are simply not added.  No halo is loaded at all.
//
// Applies to T = 8 (conv2 stage and the stem of R(2+1)D at 8-frame clips, network.py:22-27) and
// cout_pad = 64; the weights stay in shared memory for the life of the CTA.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "conv_tc.cuh"

namespace rnb {

constexpr int kTsumT = 8;       // frames per clip = frames per tile
constexpr int kTsumPos = 16;    // positions per tile (16 x 8 = 128 accumulator rows)
constexpr int kTsumCout = 64;   // output channels (padded)
constexpr int kTsumN = 3 * kTsumCout;
constexpr int kTsumMaxKb = 4;

struct ConvTsumParams {
  // activation viewed as (C, T, HW, B, 1): box (64, 8, 16, 1, 1), 128-byte swizzle
  alignas(64) CUtensorMap a_map;
  alignas(64) CUtensorMap a_map_tail;  // box (tail_ch, 8, 16, 1, 1), tail_ch * 2-byte swizzle
  alignas(64) CUtensorMap b_map;       // packed weights (K_total = 3 * cin_pad, 64), box (64, 64)
  alignas(64) CUtensorMap b_map_tail;  // box (tail_ch, 64)
  alignas(64) CUtensorMap r_map;       // residual viewed like the activation, box (64, 8, 16, 1, 1)
  const float* shift;
  const __nv_bfloat16* residual;  // (B, 8, HW, 64) or null
  __nv_bfloat16* y;               // (B, 8, HW, 64)
  int nkb;       // full 64-channel blocks of the input
  int tail_ch;   // 0, 16 or 32 trailing channels
  int cin_pad;
  int HW;        // positions per frame
  int tiles_hw;  // ceil(HW / 16)
  int B;
  int p_bufs;    // activation tiles in flight
  int r_bufs;    // residual tiles in flight (0 without residual)
  int relu;
  int reverse;   // clips last to first (see ConvPatchParams::reverse)
#ifdef RNB_PATCH_COUNTERS
  // development aid (make test_conv_cnt): 1 = every tile loads tile 0 (input from L2), 2 = no
  // output stores, 4 = no epilogue arithmetic, 8 = no MMAs
  int dbg;
#endif
};

inline int conv_tsum_smem_bytes(int nkb, int tail_ch, int p_bufs, int r_bufs) {
  const int row_bytes = nkb * 128 + tail_ch * 2;
  return 1024 /*align*/ + 1024 /*control + shift table*/ + kTsumN * row_bytes +
         p_bufs * 128 * row_bytes + r_bufs * 16384;
}

cudaError_t launch_conv_tsum(const ConvTsumParams& p, int num_sms, cudaStream_t stream);

}  // namespace rnb
