// Launchers of the non-tensor-core kernels: fp32 direct convolution (parity mode + on-device
// cross-check of the tcgen05 path) and the bandwidth-bound kernels around the convolutions.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace rnb {

// ---- geometry of the stem input as the tcgen05 path wants it --------------------------------
// "Unfolded double rows": for every (b, t, h2, wo) the 8 input columns 2*wo-3 .. 2*wo+4 times 4
// channel slots (rgb + one zero) of input row 2*h2, then the same of input row 2*h2+1: 2 x 32 bf16
// = 128 bytes.  The 7x7 stride-2 stem then is a stride-1 convolution over (h2, wo) with four
// vertical taps (row pairs ho-2 .. ho+1) of K = 64 each (window column 8, channel slot 4 and the
// upper row of the first pair carry zero weights) -- a shape the patch kernel runs.
constexpr int kStemUnfold = 32;
constexpr int kClipT = 8, kClipH = 112, kClipW = 112, kClipC = 3;

// ---- fp32 direct Conv3d, channels-last -------------------------------------------------------
struct ConvRefParams {
  const float* x;         // (B, Ti, Hi, Wi, Cin)
  const float* w;         // (kt, kh, kw, Cin, Cout)
  const float* scale;     // (Cout)
  const float* shift;     // (Cout)
  const float* residual;  // (B, To, Ho, Wo, Cout) or nullptr
  float* y;               // (B, To, Ho, Wo, Cout)
  int B, Ti, Hi, Wi, Cin;
  int To, Ho, Wo, Cout;
  int kt, kh, kw, st, sh, sw, pt, ph, pw;
  int relu;
};
cudaError_t launch_conv_ref(const ConvRefParams& p, cudaStream_t stream);

// ---- preprocessing: uint8 (B, 8, 3, 112, 112) loader frames ------------------------------------
// value = float(u8) * scale[c] + bias[c]; scale/bias are host arrays of 3 (identity = reference)
cudaError_t launch_pre_u8_to_stem(const uint8_t* in, int batch, const float* scale3,
                                  const float* bias3, __nv_bfloat16* out, cudaStream_t s);
cudaError_t launch_pre_u8_to_f32_ncdhw(const uint8_t* in, int batch, const float* scale3,
                                       const float* bias3, float* out, cudaStream_t s);
cudaError_t launch_pre_u8_to_f32_ndhwc(const uint8_t* in, int batch, const float* scale3,
                                       const float* bias3, float* out, cudaStream_t s);
cudaError_t launch_pre_u8_to_bf16_ndhwc(const uint8_t* in, int batch, const float* scale3,
                                        const float* bias3, __nv_bfloat16* out, int c_pad,
                                        cudaStream_t s);
// the same frames as the space-to-depth stem input (B, 8, 56, 56, 16) of the fused stem kernel
cudaError_t launch_pre_u8_to_s2d(const uint8_t* in, int batch, const float* scale3,
                                 const float* bias3, __nv_bfloat16* out, cudaStream_t s);
cudaError_t launch_f32_ncdhw_to_s2d(const float* in, int batch, __nv_bfloat16* out, cudaStream_t s);
// fp32 (B, 3, 8, 112, 112) (what the reference hands to the runner step) -> unfolded stem rows
cudaError_t launch_f32_ncdhw_to_stem(const float* in, int batch, __nv_bfloat16* out, cudaStream_t s);

// ---- layout changes at layer-range boundaries --------------------------------------------------
// (B, C, S) fp32  <->  (B, S, c_pad) bf16 / fp32, S = T*H*W; padded channels are written as zero
cudaError_t launch_ncdhw_f32_to_ndhwc_bf16(const float* in, int B, int C, int S, int c_pad,
                                           __nv_bfloat16* out, cudaStream_t s);
cudaError_t launch_ndhwc_bf16_to_ncdhw_f32(const __nv_bfloat16* in, int B, int C, int S, int c_pad,
                                           float* out, cudaStream_t s);
cudaError_t launch_ncdhw_f32_to_ndhwc_f32(const float* in, int B, int C, int S, float* out,
                                          cudaStream_t s);
cudaError_t launch_ndhwc_f32_to_ncdhw_f32(const float* in, int B, int C, int S, float* out,
                                          cudaStream_t s);

// ---- global average pool over S positions + Linear(C -> num_classes) ---------------------------
cudaError_t launch_pool_fc_bf16(const __nv_bfloat16* x, int B, int S, int C, const float* w,
                                const float* bias, int num_classes, float* logits, cudaStream_t s);
cudaError_t launch_pool_fc_f32(const float* x, int B, int S, int C, const float* w,
                               const float* bias, int num_classes, float* logits, cudaStream_t s);

// Same plus the per-video reduction (sum over the B clips, arg-max) in one launch.  `scratch` is a
// device buffer of pool_fc_classify_scratch_bytes(B, num_classes) bytes, 256-byte aligned; its
// contents on entry do not matter.  logits / sum / cls may each be nullptr.
size_t pool_fc_classify_scratch_bytes(int batch, int num_classes);
cudaError_t launch_pool_fc_classify_bf16(const __nv_bfloat16* x, int B, int S, int C, const float* w,
                                         const float* bias, int num_classes, float* logits,
                                         float* sum, int* cls, void* scratch, cudaStream_t s);

// Variant for callers that cannot allocate device memory: the scratch lives in a file-scope
// __device__ array inside kernels.cu (one copy per GPU that loads the library, kept until exit).
constexpr int kStaticClassifyMaxBatch = 64;
constexpr int kStaticClassifyMaxClasses = 1024;
constexpr size_t kStaticClassifyScratchBytes =
    256 + (size_t)kStaticClassifyMaxBatch * kStaticClassifyMaxClasses * sizeof(float);
// Resolves (once per device, for the CURRENT device) the device address of that array; returns
// cudaSuccess and the pointer, or the error cudaGetSymbolAddress / cudaGetDevice reported.
cudaError_t static_classify_scratch(void** ptr);
cudaError_t launch_pool_fc_classify_bf16_static(const __nv_bfloat16* x, int B, int S, int C,
                                                const float* w, const float* bias, int num_classes,
                                                float* logits, float* sum, int* cls, cudaStream_t s);

// ---- batching / aggregation ----------------------------------------------------------------------
constexpr int kMaxGatherSrcs = 64;
cudaError_t launch_batch_gather(const void* const* srcs, const int* rows, int k, void* dst,
                                size_t row_bytes, cudaStream_t s);
cudaError_t launch_logits_sum_argmax(const float* logits, int rows, int classes, float* sum,
                                     int* cls, cudaStream_t s);

}  // namespace rnb
