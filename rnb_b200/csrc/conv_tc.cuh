// Parameters of the tcgen05 implicit-GEMM convolution (shared by host planner and kernel).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace rnb {

constexpr int kMaxTaps = 32;
constexpr int kTileM = 128;          // output positions per tile = TMEM lanes
constexpr int kMaxBlockN = 256;      // accumulator columns per tile
constexpr int kTmemCols = 512;       // two accumulator buffers
constexpr int kConvThreads = 320;    // warp0 TMA, warp1 MMA, warps2-9 epilogue
constexpr int kEpiThreads = 256;
constexpr int kSmemLimit = 232448;   // 227 KB
constexpr int kMaxSplitK = 8;        // portable cluster size

// One output tile is a box of 128 output positions (bw x bh x bt x bb, powers of two) by up to
// block_n output channels.  K runs over `num_taps` filter taps times cin_pad input channels; tap i
// reads the activation view a_map[tap_map[i]] shifted by (tap_dw, tap_dh, tap_dt) positions.
struct ConvTcParams {
  alignas(64) CUtensorMap a_map[4];  // activation views, dims (C, W, H, T, B), bf16
  alignas(64) CUtensorMap b_map;     // packed weights, dims (K_total, N_rows), bf16, K-major
  __nv_bfloat16* y;                  // output (B, To, Ho, Wo, cout_pad), written by plain stores
  const float* shift;                // per output channel (BatchNorm scale is folded into the weights)
  const __nv_bfloat16* residual;     // same layout as the output, or nullptr
  int num_taps;
  int cin_pad;      // input channels per tap (multiple of 16)
  int bk;           // K elements per pipeline stage: 64 (128-byte swizzle) or 32 (64-byte swizzle)
  int kb_per_tap;   // ceil(cin_pad / bk)
  int lbw, lbh, lbt;  // log2 of the position box; bb = 128 >> (lbw + lbh + lbt)
  int tiles_w, tiles_h, tiles_t, tiles_b;
  int n_tiles, block_n;
  int Wo, Ho, To, B;  // output extent
  int cout_pad;       // stored output channels (multiple of 16)
  int relu;
  int num_stages;
  // thread-block cluster of cm x cn CTAs: the cm CTAs along M work on the same channel tile and
  // each fetch 1/cm of the weight tile, multicast to all of them; the cn CTAs along N work on the
  // same position tile and each fetch 1/cn of the activation tile (split along box dimension
  // a_slice_dim: 0 = w, 1 = h, 2 = t, 3 = b).  cm = cn = 1: no cluster.
  int cm, cn;
  // split-K variant (conv_splitk.cu): a cluster of ks CTAs shares one output tile, CTA r takes the
  // r-th contiguous range of the num_taps * kb_per_tap K blocks.  0 or 1: not split.
  int ks;
  // CTA-pair variant (conv_pair.cu): two CTAs on neighbouring position tiles share one channel tile,
  // each loading half of it, one tcgen05.mma.cta_group::2 (M = 256).  0: not paired.
  int pair;
  // Temporal-patch mode of the CTA-pair kernel (3x1x1, stride 1): a pipeline stage holds, for one
  // 64-channel block, the tile's frames plus one halo frame on either side (tp_rows rows, loaded
  // ONCE) and the weight tiles of all three taps; tap dt reads the rows from dt * tp_shift on.
  // The spatial extent is flattened (Wo = H*W, Ho = 1).  0: off.
  int tpatch, tp_rows, tp_shift;
  int m_tiles;
  int a_slice_dim, a_slice_step;
  int8_t tap_map[kMaxTaps];
  int8_t tap_dw[kMaxTaps];
  int8_t tap_dh[kMaxTaps];
  int8_t tap_dt[kMaxTaps];
  int16_t tap_coff[kMaxTaps];  // channel offset inside the activation view
  int32_t tap_koff[kMaxTaps];  // K offset inside the packed weights
  // K extent of each tap: cin_pad channels in kb_per_tap blocks for the filter taps; a fused 1x1x1
  // shortcut convolution is one more tap with its own (smaller) channel count and tensor view
  int16_t tap_cin[kMaxTaps];
  int8_t tap_nkb[kMaxTaps];
};

inline int conv_tc_stage_bytes(int bk, int block_n) { return (kTileM + block_n) * bk * 2; }

// Dynamic shared memory the kernel needs for a given configuration.
inline int conv_tc_smem_bytes(int bk, int block_n, int stages) {
  return 1024 /*alignment slack*/ + stages * conv_tc_stage_bytes(bk, block_n) + 256 /*barriers*/ +
         kMaxBlockN * 4 /*shift table*/;
}

cudaError_t launch_conv_tc(const ConvTcParams& p, int num_sms, cudaStream_t stream);

// L2 persistence of a launch's OUTPUT tensor: every kernel of the forward marks what it writes as
// "persisting" through a per-launch access-policy window, so that the next kernel -- its only
// consumer -- finds as much of it as fits the carve-out still in L2 instead of re-reading it from
// HBM (measured +1.5-3.5 % on the R18 batch-16 step).  The carve-out is the device maximum
// (cudaLimitPersistingL2CacheSize; RNB_B200_L2_PERSIST_MB caps it, 0 = off) while the largest
// tensor of the forward mostly fits it, and zero otherwise: l2_persist_configure(), called before
// the launches of a forward, switches it when the batch size changes regime.
// add_l2_output_window appends the attribute to attr[n] and returns the new count.
// cudaLimitPersistingL2CacheSize is a DEVICE-WIDE setting shared with every other stream and
// library in the context: l2_persist_acquire / _release count the model handles alive on the
// current device, and the last release puts back the limit found before the first change.
void l2_persist_configure(size_t largest_tensor_bytes);
void l2_persist_acquire();
void l2_persist_release();
int add_l2_output_window(cudaLaunchAttribute* attr, int n, const void* ptr, size_t bytes);

// Split-K variant: the fp32 staging area of the reduction (block_n x 128) aliases the operand ring.
inline int conv_splitk_smem_bytes(int bk, int block_n, int stages) {
  const int ring = stages * conv_tc_stage_bytes(bk, block_n);
  const int staging = block_n * kTileM * 4;
  return 1024 + (ring > staging ? ring : staging) + 256 + kMaxBlockN * 4;
}
cudaError_t launch_conv_splitk(const ConvTcParams& p, int num_sms, cudaStream_t stream);

// CTA-pair variant: a stage holds 128 activation rows and HALF of the weight tile.
inline int conv_pair_stage_bytes(int bk, int block_n) { return (kTileM + block_n / 2) * bk * 2; }
inline int conv_pair_smem_bytes(int bk, int block_n, int stages) {
  return 1024 + stages * conv_pair_stage_bytes(bk, block_n) + 256 + kMaxBlockN * 4;
}
inline int conv_tpatch_stage_bytes(int tp_rows, int block_n) {
  return tp_rows * 128 + 3 * (block_n / 2) * 128;
}
cudaError_t launch_conv_pair(const ConvTcParams& p, int num_sms, cudaStream_t stream);

}  // namespace rnb
