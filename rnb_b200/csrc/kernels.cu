// Bandwidth-bound kernels around the convolutions, plus the fp32 direct convolution.
// All of them are HBM- or latency-bound byte movers: coalesced, vectorised, no tensor cores.
#include "kernels.cuh"

#include <stdio.h>

#include <atomic>
#include <utility>

#include "conv_tc.cuh"
#include "ptx.cuh"

namespace rnb {

namespace {
// Launch with programmatic dependent launch allowed: the kernel may become resident while the
// previous kernel in the stream drains (its prologue overlaps that tail; measured 0.9 us instead
// of 4.1 us between kernels on a B200) and must execute griddepcontrol.wait before it touches
// anything the previous kernel reads or writes.
// `win_ptr` / `win_bytes`: the launch's output tensor for the L2 persistence window (conv_tc.cuh),
// or nullptr / 0.
template <typename... KArgs, typename... Args>
cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                       const void* win_ptr, size_t win_bytes, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = add_l2_output_window(attr, 1, win_ptr, win_bytes);
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}
}  // namespace


namespace {

constexpr int kThreads = 256;

// arg-max order of numpy (reference models/r2p1d/model.py:259,278): NaN counts as the maximum, and
// among equal candidates (or several NaNs) the lowest index wins.  True when (v, i) beats (bv, bi).
__device__ __forceinline__ bool argmax_takes(float v, int i, float bv, int bi) {
  const bool vn = v != v, bn = bv != bv;
  if (vn != bn) return vn;
  if (vn) return i < bi;
  return v > bv || (v == bv && i < bi);
}

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
// Unfolded stem rows.  A CTA stages kStemRows image rows (3 channels x 112 pixels each, read with
// coalesced 4-pixel loads) in shared memory with a 4-pixel zero margin on both sides, then every
// thread assembles 16-byte output chunks (2 input columns x 4 channel slots) and stores them
// coalesced: 336 bytes in, 3584 bytes out per image row.
constexpr int kStemRows = 8;
constexpr int kStemPitch = kClipW + 8;  // 4 zero pixels | 112 pixels | 4 zero pixels

template <typename In>
__global__ void __launch_bounds__(256)
to_stem_kernel(const In* __restrict__ in, int total_rows, Affine3 aff, uint4* __restrict__ out,
               bool ncdhw) {
  __shared__ float px[kStemRows][kClipC][kStemPitch];
  const int row0 = blockIdx.x * kStemRows;
  ptx::grid_dep_launch_dependents();
  for (int i = threadIdx.x; i < kStemRows * kClipC * 8; i += blockDim.x) {
    const int r = i / (kClipC * 8), c = (i / 8) % kClipC, m = i % 8;
    px[r][c][m < 4 ? m : kClipW + m] = 0.f;
  }
  // (launched with programmatic dependent launch: the previous kernel in the stream may still be
  // running, and the stem buffer this kernel overwrites may be one of its operands)
  ptx::grid_dep_wait();
  // one thread loads 4 consecutive pixels of one (row, channel)
  for (int i = threadIdx.x; i < kStemRows * kClipC * (kClipW / 4); i += blockDim.x) {
    const int q = i % (kClipW / 4);
    const int c = (i / (kClipW / 4)) % kClipC;
    const int r = i / (kClipC * (kClipW / 4));
    const int row = row0 + r;  // row = (b * T + t) * H + h
    if (row >= total_rows) continue;
    const int h = row % kClipH;
    const int bt = row / kClipH;
    const int t = bt % kClipT;
    const long long b = bt / kClipT;
    // uint8 frames are (B, T, C, H, W); fp32 tensors are (B, C, T, H, W)
    const long long base = ncdhw ? ((((b * kClipC + c) * kClipT + t) * kClipH + h) * (long long)kClipW)
                                 : ((((b * kClipT + t) * kClipC + c) * kClipH + h) * (long long)kClipW);
    float v[4];
    if (sizeof(In) == 1) {
      const uchar4 u = *reinterpret_cast<const uchar4*>(in + base + 4 * q);
      v[0] = u.x; v[1] = u.y; v[2] = u.z; v[3] = u.w;
    } else {
      const float4 f = *reinterpret_cast<const float4*>(in + base + 4 * q);
      v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) px[r][c][4 + 4 * q + j] = fmaf(v[j], aff.s[c], aff.b[c]);
  }
  __syncthreads();
  constexpr int kChunksPerRow = (kClipW / 2) * 4;  // 56 output positions x 4 chunks
  for (int i = threadIdx.x; i < kStemRows * kChunksPerRow; i += blockDim.x) {
    const int r = i / kChunksPerRow, o = i % kChunksPerRow;
    const int row = row0 + r;
    if (row >= total_rows) continue;
    const int wo = o >> 2, cp = o & 3;
    // columns 2*cp and 2*cp+1 of the 8-wide window that starts at input pixel 2*wo - 3;
    // window column 7 carries zero weights and is stored as zero
    const int w0 = 2 * wo - 3 + 2 * cp + 4;  // index into the padded row
    const bool second = (2 * cp + 1) < 7;
    uint4 q;
    q.x = pack2(px[r][0][w0], px[r][1][w0]);
    q.y = pack2(px[r][2][w0], 0.f);
    q.z = second ? pack2(px[r][0][w0 + 1], px[r][1][w0 + 1]) : 0u;
    q.w = second ? pack2(px[r][2][w0 + 1], 0.f) : 0u;
    // double-row layout: (b, t, h / 2, wo) x [row parity][4 chunks of 16 bytes]
    const long long bt = row / kClipH;
    const int h = row % kClipH;
    out[(((bt * (kClipH / 2) + (h >> 1)) * (kClipW / 2) + wo) * 2 + (h & 1)) * 4 + cp] = q;
  }
}

// Space-to-depth stem input for the fused stem kernel (conv_stem.cuh): (B, 8, 56, 56, 16) bf16,
// channel (py * 2 + px) * 4 + c = pixel (2 h2 + py, 2 w2 + px) of colour c; slot 3 is zero except
// for pixel (0, 0), where it holds the constant 1 (conv_stem.cuh).  One thread per four pixels of a
// row pair: six coalesced 4-pixel loads (three colours, two rows), all in flight together, and 64
// contiguous bytes out (two output positions x two rows); no shared memory, no barrier.
template <typename In>
__global__ void __launch_bounds__(128)
to_s2d_kernel(const In* __restrict__ in, long long total_quads, Affine3 aff, uint4* __restrict__ out,
              bool ncdhw) {
  constexpr int kQuads = kClipW / 4;  // 4-pixel groups per row
  ptx::grid_dep_launch_dependents();
  bool waited = false;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total_quads;
       i += (long long)gridDim.x * blockDim.x) {
    const int q = (int)(i % kQuads);
    const long long r2 = i / kQuads;  // (b * T + t) * 56 + h2
    const int h2 = (int)(r2 % (kClipH / 2));
    const long long bt = r2 / (kClipH / 2);
    const int t = (int)(bt % kClipT);
    const long long b = bt / kClipT;
    float v[2][kClipC][4];
#pragma unroll
    for (int py = 0; py < 2; ++py)
#pragma unroll
      for (int c = 0; c < kClipC; ++c) {
        const int h = 2 * h2 + py;
        const long long base =
            ncdhw ? ((((b * kClipC + c) * kClipT + t) * kClipH + h) * (long long)kClipW)
                  : ((((b * kClipT + t) * kClipC + c) * kClipH + h) * (long long)kClipW);
        if (sizeof(In) == 1) {
          const uchar4 u = *reinterpret_cast<const uchar4*>(in + base + 4 * q);
          v[py][c][0] = u.x; v[py][c][1] = u.y; v[py][c][2] = u.z; v[py][c][3] = u.w;
        } else {
          const float4 f = *reinterpret_cast<const float4*>(in + base + 4 * q);
          v[py][c][0] = f.x; v[py][c][1] = f.y; v[py][c][2] = f.z; v[py][c][3] = f.w;
        }
      }
    // The frames are read before the wait: no kernel of this library that lets its successor start
    // early (programmatic dependent launch) writes frame tensors, and anything else in the stream
    // has completed before this grid starts.  The output buffer may still be an operand of the
    // previous forward's kernels, so the stores come after it.
    if (!waited) {
      ptx::grid_dep_wait();
      waited = true;
    }
    uint4 o[4];
#pragma unroll
    for (int j = 0; j < 2; ++j)      // output position w2 = 2 q + j
#pragma unroll
      for (int py = 0; py < 2; ++py) {
        float a[kClipC][2];
#pragma unroll
        for (int c = 0; c < kClipC; ++c) {
          a[c][0] = fmaf(v[py][c][2 * j], aff.s[c], aff.b[c]);
          a[c][1] = fmaf(v[py][c][2 * j + 1], aff.s[c], aff.b[c]);
        }
        uint4 u;
        u.x = pack2(a[0][0], a[1][0]);
        // slot 3 of pixel (py, px) = (0, 0): the constant one that carries the BatchNorm shift
        u.y = pack2(a[2][0], py == 0 ? 1.f : 0.f);
        u.z = pack2(a[0][1], a[1][1]);
        u.w = pack2(a[2][1], 0.f);
        o[2 * j + py] = u;
      }
    uint4* dst = out + (r2 * (kClipW / 2) + 2 * q) * 2;
#pragma unroll
    for (int k = 0; k < 4; ++k) dst[k] = o[k];
  }
  if (!waited) ptx::grid_dep_wait();
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
// grid (B, class_chunks), kPoolThreads threads.  Latency is what this step costs (50 KB per clip
// and 0.8 MB of weights), so every load of a phase is in flight at once.  Pooling: the threads form
// blockDim / (C/2) position groups; a thread owns one channel pair and the positions of its group
// (S = 49: at most 13, unrolled), partial sums meet in shared memory.  Classifier: one class per
// warp and pass, lanes stride the C weights with 16-byte loads, shuffle reduction.
constexpr int kPoolThreads = 1024;
constexpr int kPoolClassesPerCta = kPoolThreads / 32;

template <typename In>
__device__ __forceinline__ void pool_fc_body(const In* __restrict__ x, int S, int C,
                                             const float* __restrict__ w,
                                             const float* __restrict__ bias, int num_classes,
                                             float* __restrict__ logits, float* __restrict__ logits2,
                                             float* pooled) {
  // pooled: C floats, then groups * C floats of partial sums (dynamic shared memory)
  const int b = blockIdx.x;
  const In* xb = x + (long long)b * S * C;
  const float inv = 1.f / (float)S;
  const int pairs = C / 2;
  const int groups = max(1, (int)blockDim.x / pairs);
  float* part = pooled + C;
  for (int i = threadIdx.x; i < groups * pairs; i += blockDim.x) {
    const int g = i / pairs, c = 2 * (i - g * pairs);
    float a0 = 0.f, a1 = 0.f;
#pragma unroll 13
    for (int s = g; s < S; s += groups) {
      if (sizeof(In) == 2) {
        const uint32_t u = *reinterpret_cast<const uint32_t*>(xb + (long long)s * C + c);
        a0 += __uint_as_float(u << 16);
        a1 += __uint_as_float(u & 0xffff0000u);
      } else {
        const float2 f = *reinterpret_cast<const float2*>(xb + (long long)s * C + c);
        a0 += f.x;
        a1 += f.y;
      }
    }
    part[g * C + c] = a0;
    part[g * C + c + 1] = a1;
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float a = 0.f;
    for (int g = 0; g < groups; ++g) a += part[g * C + c];  // fixed order: deterministic
    pooled[c] = a * inv;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int per_chunk = (num_classes + gridDim.y - 1) / gridDim.y;
  const int j0 = blockIdx.y * per_chunk;
  const int j1 = min(num_classes, j0 + per_chunk);
  for (int j = j0 + warp; j < j1; j += nwarps) {
    const float4* wj = reinterpret_cast<const float4*>(w + (long long)j * C);
    float acc = 0.f;
#pragma unroll 4
    for (int c4 = lane; c4 < C / 4; c4 += 32) {
      const float4 wv = __ldg(wj + c4);
      const float4 pv = *reinterpret_cast<const float4*>(pooled + 4 * c4);
      acc = fmaf(pv.x, wv.x, acc);
      acc = fmaf(pv.y, wv.y, acc);
      acc = fmaf(pv.z, wv.z, acc);
      acc = fmaf(pv.w, wv.w, acc);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
    if (lane == 0) {
      const float v = acc + bias[j];
      if (logits) logits[(long long)b * num_classes + j] = v;
      if (logits2) logits2[(long long)b * num_classes + j] = v;
    }
  }
}

// shared memory of the two kernels below: pooled values + per-group partial sums
static inline size_t pool_fc_smem_bytes(int C) {
  const int groups = kPoolThreads / (C / 2) > 0 ? kPoolThreads / (C / 2) : 1;
  return (size_t)(1 + groups) * C * sizeof(float);
}
static inline int pool_fc_chunks(int num_classes) {
  return (num_classes + kPoolClassesPerCta - 1) / kPoolClassesPerCta;
}

template <typename In>
__global__ void __launch_bounds__(kPoolThreads)
pool_fc_kernel(const In* __restrict__ x, int S, int C, const float* __restrict__ w,
               const float* __restrict__ bias, int num_classes, float* __restrict__ logits) {
  extern __shared__ float pooled[];
  ptx::grid_dep_launch_dependents();
  ptx::grid_dep_wait();  // (programmatic dependent launch: x is the previous kernel's output)
  pool_fc_body<In>(x, S, C, w, bias, num_classes, logits, nullptr, pooled);
}

// pool + classifier + per-video reduction in one launch.  Every CTA writes its slice of the per-clip
// logits to `staged` (scratch) and takes a ticket from `counter` (scratch, zero on entry); the CTA
// that draws the last ticket sums the clips in a fixed order (deterministic, unlike atomics on the
// sums), picks the arg-max with numpy's lowest-index tie rule, and re-arms the counter.
template <typename In>
__global__ void __launch_bounds__(kPoolThreads)
pool_fc_classify_kernel(const In* __restrict__ x, int S, int C, const float* __restrict__ w,
                        const float* __restrict__ bias, int num_classes,
                        float* __restrict__ logits, float* __restrict__ sum, int* __restrict__ cls,
                        float* __restrict__ staged, unsigned int* __restrict__ counter) {
  extern __shared__ float pooled[];
  __shared__ unsigned int ticket;
  __shared__ float best_v[kPoolThreads / 32];
  __shared__ int best_i[kPoolThreads / 32];
  pool_fc_body<In>(x, S, C, w, bias, num_classes, staged, logits, pooled);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) ticket = atomicAdd(counter, 1u);
  __syncthreads();
  if (ticket != gridDim.x * gridDim.y - 1) return;
  __threadfence();
  const int B = gridDim.x;
  float bv = -INFINITY;
  int bi = 0x7fffffff;
  for (int j = threadIdx.x; j < num_classes; j += blockDim.x) {
    float acc = 0.f;
    for (int b = 0; b < B; ++b) acc += __ldcg(staged + (long long)b * num_classes + j);
    if (sum) sum[j] = acc;
    if (argmax_takes(acc, j, bv, bi)) {
      bv = acc;
      bi = j;
    }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, bv, off);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, off);
    if (argmax_takes(ov, oi, bv, bi)) {
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
  if (threadIdx.x == 0) {
    for (int k = 1; k < (int)(blockDim.x >> 5); ++k)
      if (argmax_takes(best_v[k], best_i[k], bv, bi)) {
        bv = best_v[k];
        bi = best_i[k];
      }
    if (cls) *cls = bi;
    *counter = 0u;  // ready for the next launch on this scratch buffer
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
    if (argmax_takes(acc, j, bv, bi)) {
      bv = acc;
      bi = j;
    }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, bv, off);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, off);
    if (argmax_takes(ov, oi, bv, bi)) {
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
      if (argmax_takes(ov, oi, bv, bi)) {
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
  const int rows = batch * kClipT * kClipH;
  if (rows == 0) return cudaSuccess;
  return launch_pdl(to_stem_kernel<uint8_t>, dim3((rows + kStemRows - 1) / kStemRows), dim3(256), 0,
                    s, out, (size_t)rows * (kClipW / 2) * kStemUnfold * 2, in, rows, make_affine(scale3, bias3), reinterpret_cast<uint4*>(out), false);
}

cudaError_t launch_f32_ncdhw_to_stem(const float* in, int batch, __nv_bfloat16* out,
                                     cudaStream_t s) {
  const int rows = batch * kClipT * kClipH;
  if (rows == 0) return cudaSuccess;
  return launch_pdl(to_stem_kernel<float>, dim3((rows + kStemRows - 1) / kStemRows), dim3(256), 0, s,
                    nullptr, 0, in, rows, make_affine(nullptr, nullptr), reinterpret_cast<uint4*>(out), true);
}

// grid of the space-to-depth staging kernel: one thread per 4-pixel group of a row pair
static inline int s2d_grid(long long quads) { return (int)((quads + 127) / 128); }

cudaError_t launch_pre_u8_to_s2d(const uint8_t* in, int batch, const float* scale3,
                                 const float* bias3, __nv_bfloat16* out, cudaStream_t s) {
  const int rows2 = batch * kClipT * (kClipH / 2);
  if (rows2 == 0) return cudaSuccess;
  const long long quads = (long long)rows2 * (kClipW / 4);
  return launch_pdl(to_s2d_kernel<uint8_t>, dim3(s2d_grid(quads)), dim3(128), 0, s, out,
                    (size_t)rows2 * (kClipW / 2) * 32, in, quads, make_affine(scale3, bias3),
                    reinterpret_cast<uint4*>(out), false);
}

cudaError_t launch_f32_ncdhw_to_s2d(const float* in, int batch, __nv_bfloat16* out, cudaStream_t s) {
  const int rows2 = batch * kClipT * (kClipH / 2);
  if (rows2 == 0) return cudaSuccess;
  const long long quads = (long long)rows2 * (kClipW / 4);
  return launch_pdl(to_s2d_kernel<float>, dim3(s2d_grid(quads)), dim3(128), 0, s, nullptr, 0, in, quads,
                    make_affine(nullptr, nullptr), reinterpret_cast<uint4*>(out), true);
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
  return launch_pdl(pool_fc_kernel<__nv_bfloat16>, dim3(B, pool_fc_chunks(num_classes)),
                    dim3(kPoolThreads), pool_fc_smem_bytes(C), s, nullptr, 0, x, S, C, w, bias,
                    num_classes, logits);
}
cudaError_t launch_pool_fc_f32(const float* x, int B, int S, int C, const float* w,
                               const float* bias, int num_classes, float* logits, cudaStream_t s) {
  if (B == 0) return cudaSuccess;
  return launch_pdl(pool_fc_kernel<float>, dim3(B, pool_fc_chunks(num_classes)), dim3(kPoolThreads),
                    pool_fc_smem_bytes(C), s, nullptr, 0, x, S, C, w, bias, num_classes, logits);
}

size_t pool_fc_classify_scratch_bytes(int batch, int num_classes) {
  return 256 + (size_t)batch * (size_t)num_classes * sizeof(float);
}

cudaError_t launch_pool_fc_classify_bf16(const __nv_bfloat16* x, int B, int S, int C, const float* w,
                                         const float* bias, int num_classes, float* logits,
                                         float* sum, int* cls, void* scratch, cudaStream_t s) {
  if (B == 0) return cudaSuccess;
  // the counter must be zero on entry whatever the previous user of the scratch left behind
  cudaError_t e = cudaMemsetAsync(scratch, 0, 256, s);
  if (e != cudaSuccess) return e;
  unsigned int* counter = static_cast<unsigned int*>(scratch);
  float* staged = reinterpret_cast<float*>(static_cast<char*>(scratch) + 256);
  pool_fc_classify_kernel<__nv_bfloat16><<<dim3(B, pool_fc_chunks(num_classes)), kPoolThreads,
                                           pool_fc_smem_bytes(C), s>>>(
      x, S, C, w, bias, num_classes, logits, sum, cls, staged, counter);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------ static scratch
// Scratch of the no-allocation classifier entry point.  In HOST code this name is only a handle for
// the runtime, not a device address: passing it to a kernel would fault (or silently touch host
// memory on systems where the GPU can reach it).  The device address comes from
// cudaGetSymbolAddress, per device, below.
static __device__ __align__(256) unsigned char g_static_classify_scratch[kStaticClassifyScratchBytes];

cudaError_t static_classify_scratch(void** ptr) {
  constexpr int kMaxDevices = 64;
  // (atomic: two host threads making the first call on one device both resolve the same address)
  static std::atomic<void*> resolved[kMaxDevices] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
  if (resolved[dev].load() == nullptr) {
    // set-up, once per device, with that device current: every GPU has its own copy of the array
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_classify_scratch);
    if (e != cudaSuccess) return e;
    resolved[dev].store(p);
  }
  *ptr = resolved[dev].load();
  return cudaSuccess;
}

cudaError_t launch_pool_fc_classify_bf16_static(const __nv_bfloat16* x, int B, int S, int C,
                                                const float* w, const float* bias, int num_classes,
                                                float* logits, float* sum, int* cls, cudaStream_t s) {
  void* scratch = nullptr;
  cudaError_t e = static_classify_scratch(&scratch);
  if (e != cudaSuccess) return e;
  // same launcher as the caller-scratch entry point; it zeroes the arrival counter first, so
  // whatever the previous call left in the shared array does not matter
  return launch_pool_fc_classify_bf16(x, B, S, C, w, bias, num_classes, logits, sum, cls, scratch, s);
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
