// Split-K variant of the tap kernel (conv_tc.cu) for the small-M, weight-heavy layers (conv4/conv5
// at small batch): too few output tiles to occupy 148 SMs, and per tile more operand bytes than the
// SM's inbound TMA rate can deliver behind the tensor pipe.
//
// A thread-block cluster of ks CTAs works on ONE output tile (128 positions x up to 256 channels);
// CTA r accumulates the r-th contiguous range of the K blocks (filter tap x 64 input channels) in
// TMEM.  The partial accumulators are then reduced through distributed shared memory:
//
//   every CTA   accumulator -> its own shared memory as fp32 [column][128 rows] (the operand ring is
//               idle by then and is reused as the staging area)
//   cluster barrier
//   CTA r       takes the 16-column chunks c with c % ks == r: sums the ks partials of each
//               (ld.shared::cluster), adds the folded BatchNorm shift (+ residual), ReLU, packs bf16
//               and stores 32 bytes per output position, exactly like the other kernels' epilogue
//   cluster barrier (the ring may be refilled)
//
// The summation order over the K ranges is fixed (rank 0, 1, ...), so results are deterministic.
// Roles and pipeline are those of conv_tc.cu: warp 0 TMA producer, warp 1 MMA issuer, warps 2-9
// epilogue; all 320 threads take part in the two cluster barriers of every work item.
#include <atomic>
#include <stdio.h>

#include "conv_tc.cuh"
#include "ptx.cuh"

namespace rnb {

namespace {

struct SkTile {
  int w0, h0, t0, b0, n0, ncols;
};

__device__ __forceinline__ SkTile sk_decode(const ConvTcParams& p, int item) {
  SkTile c;
  const int nt = item % p.n_tiles;
  int m = item / p.n_tiles;
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

__device__ __forceinline__ float ld_shared_cluster_f32(uint32_t cluster_addr) {
  float v;
  asm volatile("ld.shared::cluster.f32 %0, [%1];\n" : "=f"(v) : "r"(cluster_addr));
  return v;
}

}  // namespace

template <bool kRes, bool kRelu>
__global__ void __launch_bounds__(kConvThreads, 1)
conv_splitk_kernel(const __grid_constant__ ConvTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  const uint32_t a_bytes = kTileM * p.bk * 2;
  const uint32_t b_bytes = p.block_n * p.bk * 2;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  // the staging area (block_n columns x 128 rows fp32) aliases the ring; the host sized the
  // dynamic shared memory for the larger of the two
  const uint32_t ring_bytes = p.num_stages * stage_bytes;
  const uint32_t staging_bytes = (uint32_t)p.block_n * kTileM * 4;
  const uint32_t bars = smem_base + (ring_bytes > staging_bytes ? ring_bytes : staging_bytes);
  const uint32_t full_bar = bars;
  const uint32_t empty_bar = bars + 8 * p.num_stages;
  const uint32_t tmem_full_bar = bars + 16 * p.num_stages;
  const uint32_t tmem_ptr_smem = tmem_full_bar + 8;
  const uint32_t shift_tab = bars + 256;  // kMaxBlockN floats
  const uint32_t staging = smem_base;

  const int ks = p.ks;
  const uint32_t crank = (uint32_t)(blockIdx.x % ks);  // provably warp-uniform (see conv_tc.cu)
  const int total_kb = p.num_taps * p.kb_per_tap;
  const int kb_per_rank = (total_kb + ks - 1) / ks;
  const int kb_begin = crank * kb_per_rank;
  const int kb_end = min(total_kb, kb_begin + kb_per_rank);  // the host guarantees kb_begin < kb_end
  const int total_items = p.m_tiles * p.n_tiles;
  const int first_item = blockIdx.x / ks;
  const int item_step = gridDim.x / ks;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&p.a_map[0]);
    ptx::prefetch_tmap(&p.b_map);
    for (int s = 0; s < p.num_stages; ++s) {
      ptx::mbar_init(full_bar + 8 * s, 1);
      ptx::mbar_init(empty_bar + 8 * s, 1);
    }
    ptx::mbar_init(tmem_full_bar, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_ptr_smem, kTmemCols);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  ptx::grid_dep_launch_dependents();
  ptx::grid_dep_wait();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];\n" : "=r"(tmem_base) : "r"(tmem_ptr_smem));
  tmem_base = __shfl_sync(0xffffffffu, tmem_base, 0);

  // epilogue geometry (warps 2-9): this thread's output position inside a tile
  const int ew = warp - 2;
  const int q = warp & 3;
  const int half = ew >> 2;
  const int row = q * 32 + lane;
  const int et = threadIdx.x - 64;
  const int lbb_shift = p.lbw + p.lbh + p.lbt;
  const int iw = row & ((1 << p.lbw) - 1);
  const int ih = (row >> p.lbw) & ((1 << p.lbh) - 1);
  const int it = (row >> (p.lbw + p.lbh)) & ((1 << p.lbt) - 1);
  const int ib = row >> lbb_shift;

  uint32_t stage = 0, phase = 0;  // ring position (producer and MMA warp each keep their own copy)
  int local = 0;
  for (int item = first_item; item < total_items; item += item_step, ++local) {
    const SkTile tc = sk_decode(p, item);
    if (warp == 0) {
      // ---------------------------------------------------------------- TMA producer (warp-wide)
      for (int i = kb_begin; i < kb_end; ++i) {
        const int tap = i / p.kb_per_tap;
        const int kb = i - tap * p.kb_per_tap;
        ptx::mbar_wait(empty_bar + 8 * stage, phase ^ 1);
        const uint32_t fb = full_bar + 8 * stage;
        const uint32_t sa = smem_base + stage * stage_bytes;
        ptx::mbar_expect_tx_elect(fb, stage_bytes);
        ptx::tma_load_5d_elect(sa, &p.a_map[p.tap_map[tap]], fb, p.tap_coff[tap] + kb * p.bk,
                               tc.w0 + p.tap_dw[tap], tc.h0 + p.tap_dh[tap], tc.t0 + p.tap_dt[tap],
                               tc.b0);
        ptx::tma_load_2d_elect(sa + a_bytes, &p.b_map, fb, p.tap_koff[tap] + kb * p.bk, tc.n0);
        if (++stage == (uint32_t)p.num_stages) {
          stage = 0;
          phase ^= 1;
        }
      }
    } else if (warp == 1) {
      // ---------------------------------------------------------------- MMA issuer (warp-wide)
      // (the accumulator is free: the second cluster barrier of the previous item is behind us)
      const uint32_t swz = p.bk * 2;
      const uint32_t idesc = ptx::make_idesc_bf16_m128((uint32_t)tc.ncols);
      ptx::tc_fence_after();
      uint32_t accumulate = 0;
      for (int i = kb_begin; i < kb_end; ++i) {
        const int kb = i % p.kb_per_tap;
        ptx::mbar_wait(full_bar + 8 * stage, phase);
        ptx::tc_fence_after();
        const uint32_t sa = smem_base + stage * stage_bytes;
        const uint64_t adesc = ptx::make_kmajor_desc(sa, swz);
        const uint64_t bdesc = ptx::make_kmajor_desc(sa + a_bytes, swz);
        const int ksteps = min(p.bk, p.cin_pad - kb * p.bk) >> 4;
        ptx::umma_bf16_elect(tmem_base, adesc, bdesc, idesc, accumulate);
        if (ksteps > 1) ptx::umma_bf16_elect(tmem_base, adesc + 2, bdesc + 2, idesc, 1);
        if (ksteps > 2) ptx::umma_bf16_elect(tmem_base, adesc + 4, bdesc + 4, idesc, 1);
        if (ksteps > 3) ptx::umma_bf16_elect(tmem_base, adesc + 6, bdesc + 6, idesc, 1);
        accumulate = 1;
        ptx::umma_commit_elect(empty_bar + 8 * stage);
        if (++stage == (uint32_t)p.num_stages) {
          stage = 0;
          phase ^= 1;
        }
      }
      ptx::umma_commit_elect(tmem_full_bar);
    } else {
      // ---------------------------------------------------------------- epilogue, part 1
      // shift table of this channel tile (the previous item's readers passed a cluster barrier)
      if (et < tc.ncols) {
        const float v = __ldg(p.shift + tc.n0 + et);
        asm volatile("st.shared.f32 [%0], %1;\n" ::"r"(shift_tab + et * 4), "f"(v) : "memory");
      }
      // my partial accumulator -> staging[column][row]; once the accumulator is complete every
      // read of the ring by the tensor core has completed too, so the ring may be overwritten
      ptx::mbar_wait(tmem_full_bar, local & 1);
      ptx::tc_fence_after();
      const uint32_t taddr_row = tmem_base + ((uint32_t)(q * 32) << 16);
      const int nchunk = tc.ncols >> 4;
      for (int c = half; c < nchunk; c += 2) {
        uint32_t v[16];
        ptx::tmem_ld16(taddr_row + c * 16, v);
        ptx::tmem_ld_wait();
        const uint32_t dst = staging + ((uint32_t)(c * 16) * kTileM + row) * 4;
#pragma unroll
        for (int j = 0; j < 16; ++j)
          asm volatile("st.shared.b32 [%0], %1;\n" ::"r"(dst + j * kTileM * 4), "r"(v[j]) : "memory");
      }
      ptx::tc_fence_before();
    }
    // every CTA's partial is staged (release / acquire at cluster scope)
    ptx::cluster_sync_all();
    if (warp >= 2) {
      // ---------------------------------------------------------------- epilogue, part 2
      const int w = tc.w0 + iw, h = tc.h0 + ih, t = tc.t0 + it, b = tc.b0 + ib;
      const bool valid = (w < p.Wo) && (h < p.Ho) && (t < p.To) && (b < p.B);
      const size_t off = ((((size_t)b * p.To + t) * p.Ho + h) * p.Wo + w) * p.cout_pad + tc.n0;
      __nv_bfloat16* out_row = p.y + off;
      const __nv_bfloat16* res_row = kRes ? p.residual + off : nullptr;
      const int nchunk = tc.ncols >> 4;
      // my chunks: c = crank + ks * j, alternating between the two warps of a lane quarter
      for (int c = (int)crank + ks * half; c < nchunk; c += 2 * ks) {
        uint32_t rwd[8];
        if (kRes) {
#pragma unroll
          for (int j = 0; j < 8; ++j) rwd[j] = 0u;
          if (valid) ptx::ld_global_nc_v8(res_row + c * 16, rwd);
        }
        float acc[16];
        const uint32_t src = staging + ((uint32_t)(c * 16) * kTileM + row) * 4;
#pragma unroll
        for (int j = 0; j < 16; ++j) acc[j] = 0.f;
        for (int r = 0; r < ks; ++r) {
          const uint32_t peer = ptx::map_to_cta(src, (uint32_t)r);
          float part[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) part[j] = ld_shared_cluster_f32(peer + j * kTileM * 4);
#pragma unroll
          for (int j = 0; j < 16; ++j) acc[j] += part[j];
        }
        uint32_t o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float2 sh;
          asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];\n"
                       : "=f"(sh.x), "=f"(sh.y)
                       : "r"(shift_tab + (c * 16 + 2 * j) * 4));
          uint64_t x = ptx::add_f32x2(ptx::pack_f32x2(acc[2 * j], acc[2 * j + 1]),
                                      ptx::pack_f32x2(sh.x, sh.y));
          if (kRes)
            x = ptx::add_f32x2(x, ptx::pack_f32x2(__uint_as_float(rwd[j] << 16),
                                                  __uint_as_float(rwd[j] & 0xffff0000u)));
          o[j] = ptx::cvt_bf16x2<kRelu>(x);
        }
        if (valid) ptx::st_global_v8(out_row + c * 16, o);
      }
      // the staging area was written through the generic proxy; the next item's TMA loads go
      // through the async proxy
      ptx::fence_proxy_async_smem();
    }
    // every CTA is done reading every staging area: the rings may be refilled
    ptx::cluster_sync_all();
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, kTmemCols);
  }
}

namespace {
template <bool kRes, bool kRelu>
cudaError_t launch_sk_variant(const ConvTcParams& p, int num_sms, int smem, cudaStream_t stream) {
  static std::atomic<bool> configured[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    e = cudaFuncSetAttribute(conv_splitk_kernel<kRes, kRelu>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit);
    if (e != cudaSuccess) return e;
    configured[dev] = true;
  }
  const int items = p.m_tiles * p.n_tiles;
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(kConvThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[3];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  attr[1].id = cudaLaunchAttributeClusterDimension;
  attr[1].val.clusterDim.x = p.ks;
  attr[1].val.clusterDim.y = 1;
  attr[1].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  // how many clusters of this size the device can hold at once (GPC boundaries strand SMs)
  static std::atomic<int> cached[64][kMaxSplitK + 1] = {};
  int max_clusters = num_sms / p.ks;
  if (dev >= 0 && dev < 64) {
    if (cached[dev][p.ks] == 0) {
      cfg.gridDim = dim3(p.ks * (num_sms / p.ks));
      int n = 0;
      e = cudaOccupancyMaxActiveClusters(&n, conv_splitk_kernel<kRes, kRelu>, &cfg);
      if (e != cudaSuccess) return e;
      cached[dev][p.ks] = n > 0 ? n : 1;
    }
    if (cached[dev][p.ks] < max_clusters) max_clusters = cached[dev][p.ks];
  }
  const int clusters = items < max_clusters ? items : max_clusters;
  cfg.gridDim = dim3(clusters * p.ks);
  cfg.numAttrs = add_l2_output_window(attr, cfg.numAttrs, p.y,
                                      (size_t)p.B * p.To * p.Ho * p.Wo * p.cout_pad * 2);
  return cudaLaunchKernelEx(&cfg, conv_splitk_kernel<kRes, kRelu>, p);
}
}  // namespace

cudaError_t launch_conv_splitk(const ConvTcParams& p, int num_sms, cudaStream_t stream) {
  if (p.ks < 2 || p.ks > kMaxSplitK || p.cm != 1 || p.cn != 1) return cudaErrorInvalidValue;
  const int total_kb = p.num_taps * p.kb_per_tap;
  const int per = (total_kb + p.ks - 1) / p.ks;
  if ((p.ks - 1) * per >= total_kb) return cudaErrorInvalidValue;  // a rank without K blocks
  const int smem = conv_splitk_smem_bytes(p.bk, p.block_n, p.num_stages);
  if (smem > kSmemLimit) return cudaErrorInvalidValue;
  const bool res = p.residual != nullptr, relu = p.relu != 0;
  if (res && relu) return launch_sk_variant<true, true>(p, num_sms, smem, stream);
  if (res) return launch_sk_variant<true, false>(p, num_sms, smem, stream);
  if (relu) return launch_sk_variant<false, true>(p, num_sms, smem, stream);
  return launch_sk_variant<false, false>(p, num_sms, smem, stream);
}

}  // namespace rnb
