// Fused stem: SpatioTemporalConv(3, 64, [3,7,7], stride [1,2,2], padding [1,3,3]) of R(2+1)D
// (reference models/r2p1d/network.py:22) as ONE kernel -- 1x7x7/2 convolution 3 -> 83, folded
// BatchNorm + ReLU, 3x1x1 convolution 83 -> 64 + bias -- with the 83-channel intermediate kept in
// shared memory.  It is the one SpatioTemporalConv of the network whose weights (48 KB + 36 KB) leave
// room for that; the two-kernel path moves 77 MB of intermediate out to HBM and back at batch 16.
//
// Input layout "space to depth": (B, 8, 56, 56, 16) bf16, channel = (py * 2 + px) * 4 + c holding
// input pixel (2 h2 + py, 2 w2 + px) of colour c (slot 3 is zero but for pixel (0, 0)).  The stride-2 7x7
// filter then is a stride-1 4x4 filter over (h2, w2) with K = 16 per tap (taps that would need
// filter row / column -1 carry zero weights).
//
// Both per-channel shifts ride in the GEMMs, so the epilogues only clamp and convert: slot 3 of
// pixel (py, px) = (0, 0) holds the constant 1 and the spatial weight row of that slot at the
// centre tap holds the folded BatchNorm shift; channel 95 of the intermediate (padding: 83 real
// channels) is written as 1 and the temporal weights of that channel at the centre tap hold the
// bias.  (The shifts are thereby rounded to bf16, like the weights.)
//
// Tile: one output row, kStemTw output columns, all 8 frames = 128 accumulator rows ordered
// (t, column) with a column pitch of 16.  The patch (4 s2d rows x 16 columns x 8 frames, fetched by
// one TMA box whose dimensions are ordered (c, w, t, h)) lies in shared memory as rows
// h * 128 + t * 16 + w, so filter tap (ah, aw) is the 128 consecutive rows from ah * 128 + aw on:
// a row offset in the UMMA descriptor, as in conv_patch.cu.  The spatial accumulator goes through
// ReLU into a second shared-memory operand of 10 frames x 16 rows (zero frames at both
// ends = the temporal padding), where temporal tap dt is the rows from dt * 16 on.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace rnb {

constexpr int kStemS2dC = 16;   // channels of the space-to-depth input
constexpr int kStemMid = 96;    // pad16(83)
constexpr int kStemOut = 64;
constexpr int kStemHW = 56;     // output (and s2d) height and width
constexpr int kStemT = 8;
constexpr int kStemTw = 12;     // output columns per tile (pitch 16: 12 + 3 halo columns fit)
constexpr int kStemTilesW = (kStemHW + kStemTw - 1) / kStemTw;  // 5

struct StemFusedParams {
  alignas(64) CUtensorMap in_map;        // s2d input, dims (16, W2, T, H2, B), box (16, 16, 8, 4, 1), 32-B swizzle
  alignas(64) CUtensorMap wsp_map;       // spatial weights (K = 256, N = 96), box (16, 96), 32-B swizzle
  alignas(64) CUtensorMap wtp_map;       // temporal weights (K = 288, N = 64), box (64, 64), 128-B swizzle
  alignas(64) CUtensorMap wtp_tail_map;  // same tensor, box (32, 64), 64-B swizzle (channels 64..95)
  __nv_bfloat16* y;                      // (B, 8, 56, 56, 64)
  int B;
};

constexpr int kStemFusedSmem = 1024 /*align*/ + 1024 /*control*/ + 16 * kStemMid * 32 +
                               3 * kStemOut * 128 + 3 * kStemOut * 64 + 4 * 16384 + 1024 +
                               2 * (160 * 128 + 160 * 64);

cudaError_t launch_conv_stem_fused(const StemFusedParams& p, int num_sms, cudaStream_t stream);

}  // namespace rnb
