// Implicit-GEMM Conv3d for channels-last bf16 activations on sm_100a.
//
//   D[pos, n] = act( scale[n] * sum_{tap, c} X[pos + tap, c] * W[n, tap, c] + shift[n] (+ R[pos, n]) )
//
// Persistent, warp-specialised:
//   warp 0      TMA producer: per K-block one 5-D box of the activation (zero-filled outside the
//               tensor, which implements the convolution padding and ragged tile edges) and one 2-D
//               box of the packed weights, into a multi-stage shared-memory ring.
//   warp 1      allocates TMEM and issues tcgen05.mma (M=128, N<=256, K=16 per instruction) into one
//               of two accumulator buffers, so the epilogue of tile i overlaps the main loop of i+1.
//   warps 2-5   epilogue: tcgen05.ld the accumulator (lane == output position), folded BatchNorm /
//               bias (scale, shift), optional residual add and ReLU, bf16 pack into a swizzled
//               staging tile, TMA store (clipped at the tensor edge).
#include <stdio.h>

#include "conv_tc.cuh"
#include "ptx.cuh"

namespace rnb {

namespace {

struct TileCoord {
  int w0, h0, t0, b0, n0, ncols;
};

__device__ __forceinline__ TileCoord decode_tile(const ConvTcParams& p, int tile) {
  TileCoord c;
  const int nt = tile % p.n_tiles;
  int m = tile / p.n_tiles;
  const int iw = m % p.tiles_w;
  m /= p.tiles_w;
  const int ih = m % p.tiles_h;
  m /= p.tiles_h;
  const int it = m % p.tiles_t;
  const int ib = m / p.tiles_t;
  const int lbb = 7 - p.lbw - p.lbh - p.lbt;
  c.w0 = iw << p.lbw;
  c.h0 = ih << p.lbh;
  c.t0 = it << p.lbt;
  c.b0 = ib << lbb;
  c.n0 = nt * p.block_n;
  c.ncols = min(p.block_n, p.cout_pad - c.n0);
  return c;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

}  // namespace

__global__ void __launch_bounds__(kConvThreads, 1)
conv_tc_kernel(const __grid_constant__ ConvTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const uint32_t a_bytes = kTileM * p.bk * 2;
  const uint32_t b_bytes = p.block_n * p.bk * 2;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  const uint32_t staging = smem_base + p.num_stages * stage_bytes;
  const uint32_t bars = staging + 2 * kStagingBytes;
  // barrier layout (8 bytes each): full[S], empty[S], tmem_full[2], tmem_empty[2], then tmem ptr
  const uint32_t full_bar = bars;
  const uint32_t empty_bar = bars + 8 * p.num_stages;
  const uint32_t tmem_full_bar = bars + 16 * p.num_stages;
  const uint32_t tmem_empty_bar = tmem_full_bar + 16;
  const uint32_t tmem_ptr_smem = tmem_empty_bar + 16;

  const int total_tiles = p.tiles_w * p.tiles_h * p.tiles_t * p.tiles_b * p.n_tiles;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&p.a_map[0]);
    ptx::prefetch_tmap(&p.b_map);
    ptx::prefetch_tmap(&p.d_map);
    for (int s = 0; s < p.num_stages; ++s) {
      ptx::mbar_init(full_bar + 8 * s, 1);
      ptx::mbar_init(empty_bar + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(tmem_full_bar + 8 * s, 1);
      ptx::mbar_init(tmem_empty_bar + 8 * s, 128);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_ptr_smem, kTmemCols);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];\n" : "=r"(tmem_base) : "r"(tmem_ptr_smem));

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const TileCoord tc = decode_tile(p, tile);
        for (int tap = 0; tap < p.num_taps; ++tap) {
          const void* amap = &p.a_map[p.tap_map[tap]];
          const int cw = tc.w0 + p.tap_dw[tap];
          const int ch = tc.h0 + p.tap_dh[tap];
          const int ct = tc.t0 + p.tap_dt[tap];
          const int coff = p.tap_coff[tap];
          const int koff = p.tap_koff[tap];
          for (int kb = 0; kb < p.kb_per_tap; ++kb) {
            ptx::mbar_wait(empty_bar + 8 * stage, phase ^ 1);
            const uint32_t fb = full_bar + 8 * stage;
            const uint32_t sa = smem_base + stage * stage_bytes;
            ptx::mbar_expect_tx(fb, stage_bytes);
            ptx::tma_load_5d(sa, amap, fb, coff + kb * p.bk, cw, ch, ct, tc.b0);
            ptx::tma_load_2d(sa + a_bytes, &p.b_map, fb, koff + kb * p.bk, tc.n0);
            if (++stage == (uint32_t)p.num_stages) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      const uint32_t swz = p.bk * 2;  // bytes per operand row == swizzle span
      uint32_t stage = 0, phase = 0;
      int local = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++local) {
        const TileCoord tc = decode_tile(p, tile);
        const uint32_t acc = local & 1;
        const uint32_t acc_phase = (local >> 1) & 1;
        const uint32_t idesc = ptx::make_idesc_bf16_m128((uint32_t)tc.ncols);
        const uint32_t tmem_d = tmem_base + acc * kMaxBlockN;
        ptx::mbar_wait(tmem_empty_bar + 8 * acc, acc_phase ^ 1);
        ptx::tc_fence_after();
        uint32_t accumulate = 0;
        for (int tap = 0; tap < p.num_taps; ++tap) {
          for (int kb = 0; kb < p.kb_per_tap; ++kb) {
            ptx::mbar_wait(full_bar + 8 * stage, phase);
            ptx::tc_fence_after();
            const uint32_t sa = smem_base + stage * stage_bytes;
            const uint64_t adesc = ptx::make_kmajor_desc(sa, swz);
            const uint64_t bdesc = ptx::make_kmajor_desc(sa + a_bytes, swz);
            const int ksteps = min(p.bk, p.cin_pad - kb * p.bk) >> 4;
            for (int k = 0; k < ksteps; ++k) {
              // advancing 16 bf16 = 32 bytes along K inside the swizzle span: +2 in the address field
              ptx::umma_bf16(tmem_d, adesc + 2 * k, bdesc + 2 * k, idesc, accumulate);
              accumulate = 1;
            }
            ptx::umma_commit(empty_bar + 8 * stage);
            if (++stage == (uint32_t)p.num_stages) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
        ptx::umma_commit(tmem_full_bar + 8 * acc);
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (128 threads)
    const int q = warp & 3;           // TMEM lane quarter this warp may read
    const int row = q * 32 + lane;    // output position inside the tile
    const bool store_thread = (threadIdx.x == 64);
    const int lbb_shift = p.lbw + p.lbh + p.lbt;
    const int iw = row & ((1 << p.lbw) - 1);
    const int ih = (row >> p.lbw) & ((1 << p.lbh) - 1);
    const int it = (row >> (p.lbw + p.lbh)) & ((1 << p.lbt) - 1);
    const int ib = row >> lbb_shift;
    int local = 0;
    uint32_t nstore = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++local) {
      const TileCoord tc = decode_tile(p, tile);
      const uint32_t acc = local & 1;
      const uint32_t acc_phase = (local >> 1) & 1;
      const int w = tc.w0 + iw, h = tc.h0 + ih, t = tc.t0 + it, b = tc.b0 + ib;
      const bool row_valid = (w < p.Wo) && (h < p.Ho) && (t < p.To) && (b < p.B);
      const __nv_bfloat16* res_row = nullptr;
      if (p.residual != nullptr && row_valid) {
        res_row = p.residual + ((((size_t)b * p.To + t) * p.Ho + h) * p.Wo + w) * p.cout_pad;
      }
      ptx::mbar_wait(tmem_full_bar + 8 * acc, acc_phase);
      ptx::tc_fence_after();
      const uint32_t taddr_row = tmem_base + ((uint32_t)(q * 32) << 16) + acc * kMaxBlockN;
      const int nsub = (tc.ncols + 63) >> 6;
      for (int sub = 0; sub < nsub; ++sub) {
        const int cols = min(64, tc.ncols - sub * 64);  // multiple of 16
        const int nchunk = cols >> 4;
        const uint32_t buf = staging + (nstore & 1) * kStagingBytes;
        // the staging buffer used two stores ago must have been read by its TMA store
        if (store_thread) ptx::tma_store_wait_read<1>();
        ptx::named_bar_sync(1, 128);
#pragma unroll 1
        for (int cchunk = 0; cchunk < nchunk; ++cchunk) {
          uint32_t v[16];
          ptx::tmem_ld16(taddr_row + sub * 64 + cchunk * 16, v);
          const int n = tc.n0 + sub * 64 + cchunk * 16;
          float sc[16], sh[16];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 s4 = __ldg(reinterpret_cast<const float4*>(p.scale + n) + j);
            const float4 h4 = __ldg(reinterpret_cast<const float4*>(p.shift + n) + j);
            sc[4 * j] = s4.x; sc[4 * j + 1] = s4.y; sc[4 * j + 2] = s4.z; sc[4 * j + 3] = s4.w;
            sh[4 * j] = h4.x; sh[4 * j + 1] = h4.y; sh[4 * j + 2] = h4.z; sh[4 * j + 3] = h4.w;
          }
          float r[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) r[j] = 0.f;
          if (res_row != nullptr) {
            const uint4 r0 = __ldg(reinterpret_cast<const uint4*>(res_row + n));
            const uint4 r1 = __ldg(reinterpret_cast<const uint4*>(res_row + n) + 1);
            const uint32_t rw[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              r[2 * j] = __uint_as_float(rw[j] << 16);
              r[2 * j + 1] = __uint_as_float(rw[j] & 0xffff0000u);
            }
          }
          ptx::tmem_ld_wait();
          float o[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            float x = fmaf(__uint_as_float(v[j]), sc[j], sh[j]) + r[j];
            o[j] = p.relu ? fmaxf(x, 0.f) : x;
          }
          // two 16-byte chunks of this row; 128-byte swizzle: chunk index XOR (row mod 8)
          const int c0 = cchunk * 2;
          const uint32_t rowaddr = buf + row * 128;
          const uint32_t a0 = rowaddr + (((c0) ^ (row & 7)) << 4);
          const uint32_t a1 = rowaddr + (((c0 + 1) ^ (row & 7)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};\n" ::"r"(a0),
                       "r"(pack_bf16x2(o[0], o[1])), "r"(pack_bf16x2(o[2], o[3])),
                       "r"(pack_bf16x2(o[4], o[5])), "r"(pack_bf16x2(o[6], o[7]))
                       : "memory");
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};\n" ::"r"(a1),
                       "r"(pack_bf16x2(o[8], o[9])), "r"(pack_bf16x2(o[10], o[11])),
                       "r"(pack_bf16x2(o[12], o[13])), "r"(pack_bf16x2(o[14], o[15]))
                       : "memory");
        }
        if (sub == nsub - 1) {
          // every TMEM read of this accumulator buffer has completed: hand it back to the MMA warp
          ptx::tc_fence_before();
          ptx::mbar_arrive(tmem_empty_bar + 8 * acc);
        }
        ptx::fence_proxy_async_smem();
        ptx::named_bar_sync(1, 128);
        if (store_thread) {
          ptx::tma_store_5d(&p.d_map, buf, tc.n0 + sub * 64, tc.w0, tc.h0, tc.t0, tc.b0);
          ptx::tma_store_commit();
        }
        ++nstore;
      }
    }
    if (store_thread) ptx::tma_store_wait_all<0>();
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, kTmemCols);
  }
}

cudaError_t launch_conv_tc(const ConvTcParams& p, int num_sms, cudaStream_t stream) {
  const int smem = conv_tc_smem_bytes(p.bk, p.block_n, p.num_stages);
  static int configured_smem = 0;
  if (smem > configured_smem) {
    cudaError_t e = cudaFuncSetAttribute(conv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kSmemLimit);
    if (e != cudaSuccess) return e;
    configured_smem = kSmemLimit;
  }
  const int total_tiles = p.tiles_w * p.tiles_h * p.tiles_t * p.tiles_b * p.n_tiles;
  const int grid = total_tiles < num_sms ? total_tiles : num_sms;
  conv_tc_kernel<<<grid, kConvThreads, smem, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace rnb
