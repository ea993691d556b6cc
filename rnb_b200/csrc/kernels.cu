// Bandwidth-bound kernels around the convolutions, plus the fp32 direct convolution.
// All of them are HBM- or latency-bound byte movers: coalesced, vectorised, no tensor cores.
#include "kernels.cuh"

namespace rnb {

namespace {

constexpr int kThreads = 256;

inline int blocks_for(long long n, int threads = kThreads) {
  long long b = (n + threads - 1) / threads;
  const long long cap = 148LL * 32;  // grid-stride beyond ~32 CTAs per SM
  return (int)(b < cap ? (b > 0 ? b : 1) : cap);
}

struct Affine3 {
  float s[3];
  float b[3];
};

inline Affine3 make_affine(const float* scale3, const float* bias3) {
  Affine3 a;
  for (int c = 0; c < 3; ++c) {
    a.s[c] = scale3 ? scale3[c] : 1.f;
    a.b[c] = bias3 ? bias3[c] : 0.f;
  }
  return a;
}

__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

// ------------------------------------------------------------------------------ fp32 direct conv
__global__ void conv_ref_kernel(const ConvRefParams p) {
  const long long total = (long long)p.B * p.To * p.Ho * p.Wo * p.Cout;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(i % p.Cout);
    long long pos = i / p.Cout;
    const int wo = (int)(pos % p.Wo);
    pos /= p.Wo;
    const int ho = (int)(pos % p.Ho);
    pos /= p.Ho;
    const int to = (int)(pos % p.To);
    const int b = (int)(pos / p.To);
    float acc = 0.f;
    for (int dt = 0; dt < p.kt; ++dt) {
      const int t = to * p.st + dt - p.pt;
      if (t < 0 || t >= p.Ti) continue;
      for (int dh = 0; dh < p.kh; ++dh) {
        const int h = ho * p.sh + dh - p.ph;
        if (h < 0 || h >= p.Hi) continue;
        for (int dw = 0; dw < p.kw; ++dw) {
          const int w = wo * p.sw + dw - p.pw;
          if (w < 0 || w >= p.Wi) continue;
          const float* xp = p.x + ((((long long)b * p.Ti + t) * p.Hi + h) * p.Wi + w) * p.Cin;
          const float* wp = p.w + ((long long)((dt * p.kh + dh) * p.kw + dw) * p.Cin) * p.Cout + n;
          int c = 0;
          for (; c + 4 <= p.Cin; c += 4) {
            acc = fmaf(xp[c], wp[(long long)c * p.Cout], acc);
            acc = fmaf(xp[c + 1], wp[(long long)(c + 1) * p.Cout], acc);
            acc = fmaf(xp[c + 2], wp[(long long)(c + 2) * p.Cout], acc);
            acc = fmaf(xp[c + 3], wp[(long long)(c + 3) * p.Cout], acc);
          }
          for (; c < p.Cin; ++c) acc = fmaf(xp[c], wp[(long long)c * p.Cout], acc);
        }
      }
    }
    float v = fmaf(acc, p.scale[n], p.shift[n]);
    if (p.residual) v += p.residual[i];
    p.y[i] = p.relu ? fmaxf(v, 0.f) : v;
  }
}

// ------------------------------------------------------------------------------ preprocessing
// One thread writes one 16-byte chunk: 2 input columns x 4 channel slots of an unfolded stem row.
template <typename In>
__global__ void to_stem_kernel(const In* __restrict__ in, long long chunks, Affine3 aff,
                               uint4* __restrict__ out, bool ncdhw) {
  for (long long o = blockIdx.x * (long long)blockDim.x + threadIdx.x; o < chunks;
       o += (long long)gridDim.x * blockDim.x) {
    const int cp = (int)(o & 3);
    long long r = o >> 2;
    const int wo = (int)(r % (kClipW / 2));
    r /= (kClipW / 2);
    const int h = (int)(r % kClipH);
    r /= kClipH;
    const int t = (int)(r % kClipT);
    const long long b = r / kClipT;
    float v[2][3];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int col = cp * 2 + j;
      const int w = 2 * wo + col - 3;
      const bool ok = (col < 7) && (w >= 0) && (w < kClipW);
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        float x = 0.f;
        if (ok) {
          // uint8 frames are (B, T, C, H, W); fp32 tensors are (B, C, T, H, W)
          const long long idx = ncdhw ? ((((b * kClipC + c) * kClipT + t) * kClipH + h) * kClipW + w)
                                      : ((((b * kClipT + t) * kClipC + c) * kClipH + h) * kClipW + w);
          x = fmaf((float)in[idx], aff.s[c], aff.b[c]);
        }
        v[j][c] = x;
      }
    }
    uint4 q;
    q.x = pack2(v[0][0], v[0][1]);
    q.y = pack2(v[0][2], 0.f);
    q.z = pack2(v[1][0], v[1][1]);
    q.w = pack2(v[1][2], 0.f);
    out[o] = q;
  }
}

// uint8 (B,T,C,H,W) -> fp32 (B,C,T,H,W); four pixels per thread
__global__ void u8_to_f32_ncdhw_kernel(const uchar4* __restrict__ in, long long quads, Affine3 aff,
                                       float4* __restrict__ out) {
  constexpr int HW4 = kClipH * kClipW / 4;
  for (long long o = blockIdx.x * (long long)blockDim.x + threadIdx.x; o < quads;
       o += (long long)gridDim.x * blockDim.x) {
    // o indexes the OUTPUT (b, c, t, hw4): writes are fully coalesced, reads are 4-byte coalesced
    const int hw = (int)(o % HW4);
    long long r = o / HW4;
    const int t = (int)(r % kClipT);
    r /= kClipT;
    const int c = (int)(r % kClipC);
    const long long b = r / kClipC;
    const uchar4 u = in[((b * kClipT + t) * kClipC + c) * HW4 + hw];
    float4 f;
    f.x = fmaf((float)u.x, aff.s[c], aff.b[c]);
    f.y = fmaf((float)u.y, aff.s[c], aff.b[c]);
    f.z = fmaf((float)u.z, aff.s[c], aff.b[c]);
    f.w = fmaf((float)u.w, aff.s[c], aff.b[c]);
    out[o] = f;
  }
}

// uint8 (B,T,C,H,W) -> channels-last (B,T,H,W,c_pad); one thread per pixel
template <typename Out>
__global__ void u8_to_ndhwc_kernel(const uint8_t* __restrict__ in, long long pixels, Affine3 aff,
                                   Out* __restrict__ out, int c_pad) {
  constexpr int HW = kClipH * kClipW;
  for (long long o = blockIdx.x * (long long)blockDim.x + threadIdx.x; o < pixels;
       o += (long long)gridDim.x * blockDim.x) {
    const int hw = (int)(o % HW);
    const long long bt = o / HW;
    Out* dst = out + o * c_pad;
    for (int c = 0; c < c_pad; ++c) {
      float x = 0.f;
      if (c < kClipC) x = fmaf((float)in[(bt * kClipC + c) * HW + hw], aff.s[c], aff.b[c]);
      dst[c] = (Out)x;
    }
  }
}

// ------------------------------------------------------------------------------ layout transposes
// (B, C, S) -> (B, S, c_pad) through a 32x32 shared-memory tile; block = (32, 8)
template <typename In, typename Out>
__global__ void bcs_to_bsc_kernel(const In* __restrict__ in, int C, int S, int c_pad,
                                  Out* __restrict__ out) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z;
  const int s0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  for (int k = threadIdx.y; k < 32; k += 8) {
    const int c = c0 + k, s = s0 + threadIdx.x;
    tile[k][threadIdx.x] = (c < C && s < S) ? (float)in[((long long)b * C + c) * S + s] : 0.f;
  }
  __syncthreads();
  for (int k = threadIdx.y; k < 32; k += 8) {
    const int s = s0 + k, c = c0 + threadIdx.x;
    if (s < S && c < c_pad) out[((long long)b * S + s) * c_pad + c] = (Out)tile[threadIdx.x][k];
  }
}

// (B, S, c_pad) -> (B, C, S)
template <typename In, typename Out>
__global__ void bsc_to_bcs_kernel(const In* __restrict__ in, int C, int S, int c_pad,
                                  Out* __restrict__ out) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z;
  const int s0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  for (int k = threadIdx.y; k < 32; k += 8) {
    const int s = s0 + k, c = c0 + threadIdx.x;
    tile[k][threadIdx.x] = (s < S && c < C) ? (float)in[((long long)b * S + s) * c_pad + c] : 0.f;
  }
  __syncthreads();
  for (int k = threadIdx.y; k < 32; k += 8) {
    const int c = c0 + k, s = s0 + threadIdx.x;
    if (c < C && s < S) out[((long long)b * C + c) * S + s] = (Out)tile[threadIdx.x][k];
  }
}

// ------------------------------------------------------------------------------ pool + FC
// grid (B, class_chunks); every CTA pools its clip (S x C, coalesced over channels) into shared
// memory, then each warp produces classes with lane-strided dot products and a shuffle reduction.
template <typename In>
__global__ void pool_fc_kernel(const In* __restrict__ x, int S, int C, const float* __restrict__ w,
                               const float* __restrict__ bias, int num_classes,
                               float* __restrict__ logits) {
  extern __shared__ float pooled[];
  const int b = blockIdx.x;
  const In* xb = x + (long long)b * S * C;
  const float inv = 1.f / (float)S;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float acc = 0.f;
    for (int s = 0; s < S; ++s) acc += (float)xb[(long long)s * C + c];
    pooled[c] = acc * inv;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int per_chunk = (num_classes + gridDim.y - 1) / gridDim.y;
  const int j0 = blockIdx.y * per_chunk;
  const int j1 = min(num_classes, j0 + per_chunk);
  for (int j = j0 + warp; j < j1; j += nwarps) {
    const float* wj = w + (long long)j * C;
    float acc = 0.f;
    for (int c = lane; c < C; c += 32) acc = fmaf(pooled[c], __ldg(wj + c), acc);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
    if (lane == 0) logits[(long long)b * num_classes + j] = acc + bias[j];
  }
}

// ------------------------------------------------------------------------------ gather / reduce
struct GatherArgs {
  const void* src[kMaxGatherSrcs];
  long long dst_off[kMaxGatherSrcs];  // bytes
  long long bytes[kMaxGatherSrcs];
};

template <typename V>
__global__ void gather_kernel(const GatherArgs a, uint8_t* __restrict__ dst) {
  const int k = blockIdx.y;
  const V* s = reinterpret_cast<const V*>(a.src[k]);
  V* d = reinterpret_cast<V*>(dst + a.dst_off[k]);
  const long long n = a.bytes[k] / (long long)sizeof(V);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x)
    d[i] = s[i];
}

// one CTA: per-class sum over rows, then arg-max with the lowest index winning ties (numpy's rule)
__global__ void sum_argmax_kernel(const float* __restrict__ logits, int rows, int classes,
                                  float* __restrict__ sum, int* __restrict__ cls) {
  __shared__ float best_v[32];
  __shared__ int best_i[32];
  float bv = -INFINITY;
  int bi = 0x7fffffff;
  for (int j = threadIdx.x; j < classes; j += blockDim.x) {
    float acc = 0.f;
    for (int r = 0; r < rows; ++r) acc += logits[(long long)r * classes + j];
    if (sum) sum[j] = acc;
    if (acc > bv || (acc == bv && j < bi)) {
      bv = acc;
      bi = j;
    }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, bv, off);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, off);
    if (ov > bv || (ov == bv && oi < bi)) {
      bv = ov;
      bi = oi;
    }
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    best_v[warp] = bv;
    best_i[warp] = bi;
  }
  __syncthreads();
  if (warp == 0) {
    const int nw = blockDim.x >> 5;
    bv = lane < nw ? best_v[lane] : -INFINITY;
    bi = lane < nw ? best_i[lane] : 0x7fffffff;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, off);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, off);
      if (ov > bv || (ov == bv && oi < bi)) {
        bv = ov;
        bi = oi;
      }
    }
    if (lane == 0 && cls) *cls = bi;
  }
}

}  // namespace

// ================================================================================ launchers
cudaError_t launch_conv_ref(const ConvRefParams& p, cudaStream_t stream) {
  const long long total = (long long)p.B * p.To * p.Ho * p.Wo * p.Cout;
  if (total == 0) return cudaSuccess;
  conv_ref_kernel<<<blocks_for(total), kThreads, 0, stream>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_pre_u8_to_stem(const uint8_t* in, int batch, const float* scale3,
                                  const float* bias3, __nv_bfloat16* out, cudaStream_t s) {
  const long long chunks = (long long)batch * kClipT * kClipH * (kClipW / 2) * 4;
  if (chunks == 0) return cudaSuccess;
  to_stem_kernel<uint8_t><<<blocks_for(chunks), kThreads, 0, s>>>(
      in, chunks, make_affine(scale3, bias3), reinterpret_cast<uint4*>(out), false);
  return cudaGetLastError();
}

cudaError_t launch_f32_ncdhw_to_stem(const float* in, int batch, __nv_bfloat16* out,
                                     cudaStream_t s) {
  const long long chunks = (long long)batch * kClipT * kClipH * (kClipW / 2) * 4;
  if (chunks == 0) return cudaSuccess;
  to_stem_kernel<float><<<blocks_for(chunks), kThreads, 0, s>>>(
      in, chunks, make_affine(nullptr, nullptr), reinterpret_cast<uint4*>(out), true);
  return cudaGetLastError();
}

cudaError_t launch_pre_u8_to_f32_ncdhw(const uint8_t* in, int batch, const float* scale3,
                                       const float* bias3, float* out, cudaStream_t s) {
  const long long quads = (long long)batch * kClipT * kClipC * kClipH * kClipW / 4;
  if (quads == 0) return cudaSuccess;
  u8_to_f32_ncdhw_kernel<<<blocks_for(quads), kThreads, 0, s>>>(
      reinterpret_cast<const uchar4*>(in), quads, make_affine(scale3, bias3),
      reinterpret_cast<float4*>(out));
  return cudaGetLastError();
}

cudaError_t launch_pre_u8_to_f32_ndhwc(const uint8_t* in, int batch, const float* scale3,
                                       const float* bias3, float* out, cudaStream_t s) {
  const long long pixels = (long long)batch * kClipT * kClipH * kClipW;
  if (pixels == 0) return cudaSuccess;
  u8_to_ndhwc_kernel<float><<<blocks_for(pixels), kThreads, 0, s>>>(
      in, pixels, make_affine(scale3, bias3), out, kClipC);
  return cudaGetLastError();
}

cudaError_t launch_pre_u8_to_bf16_ndhwc(const uint8_t* in, int batch, const float* scale3,
                                        const float* bias3, __nv_bfloat16* out, int c_pad,
                                        cudaStream_t s) {
  const long long pixels = (long long)batch * kClipT * kClipH * kClipW;
  if (pixels == 0) return cudaSuccess;
  u8_to_ndhwc_kernel<__nv_bfloat16><<<blocks_for(pixels), kThreads, 0, s>>>(
      in, pixels, make_affine(scale3, bias3), out, c_pad);
  return cudaGetLastError();
}

namespace {
template <typename In, typename Out>
cudaError_t transpose_to_cl(const In* in, int B, int C, int S, int c_pad, Out* out, cudaStream_t s) {
  if (B == 0) return cudaSuccess;
  dim3 grid((S + 31) / 32, (c_pad + 31) / 32, B), block(32, 8);
  bcs_to_bsc_kernel<In, Out><<<grid, block, 0, s>>>(in, C, S, c_pad, out);
  return cudaGetLastError();
}
template <typename In, typename Out>
cudaError_t transpose_from_cl(const In* in, int B, int C, int S, int c_pad, Out* out,
                              cudaStream_t s) {
  if (B == 0) return cudaSuccess;
  dim3 grid((S + 31) / 32, (C + 31) / 32, B), block(32, 8);
  bsc_to_bcs_kernel<In, Out><<<grid, block, 0, s>>>(in, C, S, c_pad, out);
  return cudaGetLastError();
}
}  // namespace

cudaError_t launch_ncdhw_f32_to_ndhwc_bf16(const float* in, int B, int C, int S, int c_pad,
                                           __nv_bfloat16* out, cudaStream_t s) {
  return transpose_to_cl(in, B, C, S, c_pad, out, s);
}
cudaError_t launch_ndhwc_bf16_to_ncdhw_f32(const __nv_bfloat16* in, int B, int C, int S, int c_pad,
                                           float* out, cudaStream_t s) {
  return transpose_from_cl(in, B, C, S, c_pad, out, s);
}
cudaError_t launch_ncdhw_f32_to_ndhwc_f32(const float* in, int B, int C, int S, float* out,
                                          cudaStream_t s) {
  return transpose_to_cl(in, B, C, S, C, out, s);
}
cudaError_t launch_ndhwc_f32_to_ncdhw_f32(const float* in, int B, int C, int S, float* out,
                                          cudaStream_t s) {
  return transpose_from_cl(in, B, C, S, C, out, s);
}

cudaError_t launch_pool_fc_bf16(const __nv_bfloat16* x, int B, int S, int C, const float* w,
                                const float* bias, int num_classes, float* logits,
                                cudaStream_t s) {
  if (B == 0) return cudaSuccess;
  pool_fc_kernel<__nv_bfloat16><<<dim3(B, 4), 256, C * sizeof(float), s>>>(x, S, C, w, bias,
                                                                          num_classes, logits);
  return cudaGetLastError();
}
cudaError_t launch_pool_fc_f32(const float* x, int B, int S, int C, const float* w,
                               const float* bias, int num_classes, float* logits, cudaStream_t s) {
  if (B == 0) return cudaSuccess;
  pool_fc_kernel<float><<<dim3(B, 4), 256, C * sizeof(float), s>>>(x, S, C, w, bias, num_classes,
                                                                  logits);
  return cudaGetLastError();
}

cudaError_t launch_batch_gather(const void* const* srcs, const int* rows, int k, void* dst,
                                size_t row_bytes, cudaStream_t s) {
  if (k < 0 || k > kMaxGatherSrcs) return cudaErrorInvalidValue;
  GatherArgs a;
  long long off = 0, max_bytes = 0;
  bool vec = (row_bytes % 16 == 0) && (reinterpret_cast<uintptr_t>(dst) % 16 == 0);
  for (int i = 0; i < k; ++i) {
    a.src[i] = srcs[i];
    a.dst_off[i] = off;
    a.bytes[i] = (long long)rows[i] * (long long)row_bytes;
    off += a.bytes[i];
    if (a.bytes[i] > max_bytes) max_bytes = a.bytes[i];
    vec = vec && (reinterpret_cast<uintptr_t>(srcs[i]) % 16 == 0);
  }
  for (int i = k; i < kMaxGatherSrcs; ++i) {
    a.src[i] = nullptr;
    a.dst_off[i] = 0;
    a.bytes[i] = 0;
  }
  if (k == 0 || max_bytes == 0) return cudaSuccess;
  if (vec) {
    dim3 grid(blocks_for(max_bytes / 16), k);
    gather_kernel<uint4><<<grid, kThreads, 0, s>>>(a, static_cast<uint8_t*>(dst));
  } else {
    dim3 grid(blocks_for(max_bytes), k);
    gather_kernel<uint8_t><<<grid, kThreads, 0, s>>>(a, static_cast<uint8_t*>(dst));
  }
  return cudaGetLastError();
}

cudaError_t launch_logits_sum_argmax(const float* logits, int rows, int classes, float* sum,
                                     int* cls, cudaStream_t s) {
  sum_argmax_kernel<<<1, 512, 0, s>>>(logits, rows, classes, sum, cls);
  return cudaGetLastError();
}

}  // namespace rnb
