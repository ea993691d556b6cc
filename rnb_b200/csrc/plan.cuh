// Host-side planning for the tcgen05 convolution: geometry -> tile shape, tap table, TMA tensor
// maps, packed weights.
#pragma once
#include <string>
#include <vector>

#include "conv_patch.cuh"
#include "conv_stem.cuh"
#include "conv_tc.cuh"
#include "conv_tsum.cuh"

namespace rnb {

inline int round_up(int v, int m) { return (v + m - 1) / m * m; }
inline int pad16(int c) { return round_up(c, 16); }

// One Conv3d over a channels-last activation (B, Ti, Hi, Wi, cin_pad) -> (B, To, Ho, Wo, cout_pad).
struct ConvGeom {
  int Ti = 1, Hi = 1, Wi = 1, cin = 0;
  int To = 1, Ho = 1, Wo = 1, cout = 0;
  int kt = 1, kh = 1, kw = 1;
  int st = 1, sh = 1, sw = 1;
  int pt = 0, ph = 0, pw = 0;
  // stem (7x7 stride 2 on 3 channels): the input is the "unfolded double-row" layout of
  // kernels.cuh, (B, Ti, Hi/2, Wo, 64) -- per output column the 8 input columns x 4 channel slots it
  // needs, of input rows 2*h2 and 2*h2+1 -- which turns the convolution into a stride-1 one with
  // four vertical taps (row pairs h2 = ho-2 .. ho+1) of K = 64 for the patch kernel
  bool stem_unfolded = false;

  // Fused shortcut (K-concatenation): extra_cin > 0 appends ONE tap that reads extra_cin channels
  // (from channel extra_coff on) of a second tensor (B, extra_Ti, Ho, Wo, extra_pitch) at frame
  // extra_st * t -- the 1x1x1 temporal convolution of a downsample shortcut, whose result would
  // otherwise be added as a residual.  Tap kernel and CTA-pair kernel only.
  int extra_cin = 0, extra_st = 1, extra_Ti = 0, extra_pitch = 0, extra_coff = 0;
  // Fused shortcut (N-concatenation): ncat > 0 appends ncat output channels, from column
  // pad16(cout) on, that see only the filter tap at offset zero (dt = pt, dh = ph, dw = pw) -- the
  // 1x1x1 spatial convolution of a downsample shortcut, which reads exactly the positions that tap
  // of the block's first (equally strided) convolution reads.  Tap / CTA-pair kernel only.
  int ncat = 0;
  // channels per position of the INPUT tensor in memory when it is a slice of a wider tensor
  // (0: cin_pad())
  int in_pitch = 0;

  int cin_pad() const { return stem_unfolded ? 64 : pad16(cin); }
  int cout_pad() const { return pad16(cout) + (ncat ? pad16(ncat) : 0); }
  int num_taps() const { return stem_unfolded ? 4 : kt * kh * kw; }
  int extra_cin_pad() const { return extra_cin ? pad16(extra_cin) : 0; }
  int k_total() const { return num_taps() * cin_pad() + extra_cin_pad(); }
  long long in_elems(int B) const {
    return stem_unfolded ? (long long)B * Ti * (Hi / 2) * Wo * 64
                         : (long long)B * Ti * Hi * Wi * (in_pitch ? in_pitch : cin_pad());
  }
  long long out_elems(int B) const { return (long long)B * To * Ho * Wo * cout_pad(); }
  double flops(int B) const {  // 2 x MACs, zero-pad taps counted, real channels only
    return 2.0 * B * To * Ho * Wo *
           ((double)cout * ((double)cin * kt * kh * kw + extra_cin) + (double)ncat * cin);
  }
};

// torch weight (cout, cin, kt, kh, kw) fp32 -> K-major bf16 rows (cout_pad, k_total); row n is
// multiplied by scale[n] (the folded BatchNorm scale) before rounding when `scale` is given
// (w_extra (cout, extra_cin) / scale_extra: the K-concatenated shortcut convolution's weights and
// scale; w_ncat (ncat, cin) / scale_ncat: the N-concatenated one's)
void pack_conv_weights_tc(const ConvGeom& g, const float* w, const float* scale,
                          std::vector<uint16_t>* packed, const float* w_extra = nullptr,
                          const float* scale_extra = nullptr, const float* w_ncat = nullptr,
                          const float* scale_ncat = nullptr);

// torch weight (cout, cin, kt, kh, kw) -> (kt, kh, kw, cin, cout) fp32 for the direct convolution
void pack_conv_weights_ref(const ConvGeom& g, const float* w, std::vector<float>* packed);

// Fill `out` for a batch of B clips.  All pointers are device pointers.  Returns false and sets
// *err on failure (unsupported geometry or a tensor-map encoding error).
// force_cm/force_cn > 0 pin the cluster shape (falling back to 1 x 1 when the geometry cannot be
// split that way); 0 lets the traffic model choose.
bool build_conv_tc_params(const ConvGeom& g, int B, const __nv_bfloat16* x,
                          const __nv_bfloat16* w_packed, __nv_bfloat16* y, const float* shift,
                          const __nv_bfloat16* residual, bool relu, int num_sms, ConvTcParams* out,
                          std::string* err, int force_cm = 0, int force_cn = 0, int split_k = 0,
                          int split_n = 0, int pair = 0, const __nv_bfloat16* x_extra = nullptr);
// (split_k > 1 plans the split-K variant, conv_splitk.cu: a cluster of split_k CTAs per output tile;
// split_n caps its channel tile, default = the accumulator width; pair = 1 plans the CTA-pair
// variant, conv_pair.cu, whose channel tile split_n caps as well)

// Temporal-patch mode of the CTA-pair kernel (conv_pair.cu): 3x1x1 stride-1 convolutions whose
// input is too wide for the patch kernel (conv3 / conv4 temporal).  Each 64-channel block of the
// tile's frames + halo is loaded once and serves all three taps.  n_cap caps the channel tile.
bool conv_tpatch_applicable(const ConvGeom& g);
bool build_conv_tpatch_params(const ConvGeom& g, int B, const __nv_bfloat16* x,
                              const __nv_bfloat16* w_packed, __nv_bfloat16* y, const float* shift,
                              const __nv_bfloat16* residual, bool relu, int n_cap, ConvTcParams* out,
                              std::string* err);

// Patch kernel (conv_patch.cuh): stride-1 1x3x3 and 3x1x1 convolutions; weights resident in shared
// memory when they fit next to two activation patches, streamed otherwise.  pair = 1 plans the
// CTA-pair (cta_group::2) variant, in which each CTA holds half of every weight tile.
bool conv_patch_applicable(const ConvGeom& g, int pair = 0);
// true when the patch kernel would have to stream its weights (they do not fit in shared memory);
// then the tap kernel may be faster and the model times both
bool conv_patch_streams(const ConvGeom& g, int pair = 0);
bool build_conv_patch_params(const ConvGeom& g, int B, const __nv_bfloat16* x,
                             const __nv_bfloat16* w_packed, __nv_bfloat16* y, const float* shift,
                             const __nv_bfloat16* residual, bool relu, ConvPatchParams* out,
                             std::string* err, int pair = 0);

// Temporal 3x1x1 convolution with the taps stacked along N and summed in the epilogue
// (conv_tsum.cuh): 8-frame clips, 64 output channels, weights resident.
bool conv_tsum_applicable(const ConvGeom& g);
bool build_conv_tsum_params(const ConvGeom& g, int B, const __nv_bfloat16* x,
                            const __nv_bfloat16* w_packed, __nv_bfloat16* y, const float* shift,
                            const __nv_bfloat16* residual, bool relu, ConvTsumParams* out,
                            std::string* err);

// Fused stem (conv_stem.cuh).  Spatial weights: torch (cout, 3, 1, 7, 7) fp32 -> K-major bf16 rows
// (96, 256), K = (4x4 tap of the space-to-depth filter) x 16 channel slots, rows scaled by the
// folded BatchNorm, plus the shift itself in the constant-one slot of the centre tap.  The temporal
// weights are pack_conv_weights_tc's (64 rows x 3 x 96) with the bias added in the constant-one
// channel of the centre tap (stem_fused_add_temporal_bias).
void pack_stem_fused_weights(const float* w, const float* scale, const float* shift, int cout,
                             std::vector<uint16_t>* packed);
void stem_fused_add_temporal_bias(const float* bias, int cout, std::vector<uint16_t>* packed);
bool build_stem_fused_params(int B, const __nv_bfloat16* s2d, const __nv_bfloat16* w_spatial,
                             const __nv_bfloat16* w_temporal, __nv_bfloat16* y,
                             StemFusedParams* out, std::string* err);

uint16_t f32_to_bf16_bits(float f);
float bf16_bits_to_f32(uint16_t b);

}  // namespace rnb
