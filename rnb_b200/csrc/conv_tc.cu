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
//   warps 2-9   epilogue: tcgen05.ld the accumulator (lane == output position), folded BatchNorm /
//               bias (shift; the scale lives in the weights), optional residual add and ReLU, bf16
//               pack, one 32-byte global store per 16 channels (rows outside the tensor masked).
#include <atomic>
#include <stdio.h>
#include <stdlib.h>

#include "conv_tc.cuh"
#include "epilogue.cuh"
#include "ptx.cuh"

namespace rnb {

namespace {

struct TileCoord {
  int w0, h0, t0, b0, n0, ncols;
};

// position tile m (may be >= m_tiles for the padding CTAs of a cluster: its batch origin then
// lies outside the tensor) and channel tile nt (ncols <= 0 when it lies beyond the last tile)
__device__ __forceinline__ TileCoord decode_tile(const ConvTcParams& p, int m, int nt) {
  TileCoord c;
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

// work items of a cluster: groups of cm position tiles x cn channel tiles, channel group fastest
struct ItemIter {
  int n_groups, total, rm, rn, cm, cn;
  __device__ __forceinline__ void tile(int item, int* m, int* nt) const {
    const int ng = item % n_groups;
    const int mg = item / n_groups;
    *m = mg * cm + rm;
    *nt = ng * cn + rn;
  }
};

}  // namespace

template <bool kRes, bool kRelu>
__global__ void __launch_bounds__(kConvThreads, 1)
conv_tc_kernel(const __grid_constant__ ConvTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  // shuffle-broadcast so the compiler knows these are warp-uniform (role branches, TMEM address)
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  const uint32_t a_bytes = kTileM * p.bk * 2;
  const uint32_t b_bytes = p.block_n * p.bk * 2;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  const uint32_t bars = smem_base + p.num_stages * stage_bytes;
  // barrier layout (8 bytes each): full[S], empty[S], tmem_full[2], tmem_empty[2], then tmem ptr
  const uint32_t full_bar = bars;
  const uint32_t empty_bar = bars + 8 * p.num_stages;
  const uint32_t tmem_full_bar = bars + 16 * p.num_stages;
  const uint32_t tmem_empty_bar = tmem_full_bar + 16;
  const uint32_t tmem_ptr_smem = tmem_empty_bar + 16;
  const uint32_t shift_tab = bars + 256;  // kMaxBlockN floats

  // ---- cluster geometry
  const int csize = p.cm * p.cn;
  // rank inside the (csize,1,1) cluster of a 1-D grid = blockIdx.x % csize, which -- unlike a read
  // of %cluster_ctarank -- the compiler can prove warp-uniform (keeps the MMA issue loop short)
  const uint32_t crank = csize > 1 ? (uint32_t)(blockIdx.x % csize) : 0u;
  ItemIter items;
  items.cm = p.cm;
  items.cn = p.cn;
  items.rm = crank % p.cm;
  items.rn = crank / p.cm;
  items.n_groups = (p.n_tiles + p.cn - 1) / p.cn;
  items.total = ((p.m_tiles + p.cm - 1) / p.cm) * items.n_groups;
  const int first_item = blockIdx.x / csize;
  const int item_step = gridDim.x / csize;
  uint16_t a_mask = 0, b_mask = 0;  // CTAs that need my slice of A (same rm) / of B (same rn)
  for (int j = 0; j < p.cn; ++j) a_mask |= (uint16_t)(1u << (items.rm + j * p.cm));
  for (int i = 0; i < p.cm; ++i) b_mask |= (uint16_t)(1u << (i + items.rn * p.cm));
  const uint16_t peers_mask = a_mask | b_mask;  // everyone that writes into my stages, me included
  const uint32_t a_slice_bytes = a_bytes / p.cn;
  const uint32_t b_slice_bytes = b_bytes / p.cm;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&p.a_map[0]);
    ptx::prefetch_tmap(&p.b_map);
    for (int s = 0; s < p.num_stages; ++s) {
      ptx::mbar_init(full_bar + 8 * s, 1);
      ptx::mbar_init(empty_bar + 8 * s, p.cm + p.cn - 1);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(tmem_full_bar + 8 * s, 1);
      ptx::mbar_init(tmem_empty_bar + 8 * s, kEpiThreads / 32);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_ptr_smem, kTmemCols);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  // peers multicast into my stages and arrive on my barriers: nobody may start before every CTA
  // of the cluster has initialised its barriers
  if (csize > 1) ptx::cluster_sync_all();
  ptx::tc_fence_after();
  // everything above touched only shared memory, TMEM and kernel parameters; from here on we read
  // what the previous kernel in the stream wrote and overwrite what it may still be reading
  ptx::grid_dep_launch_dependents();
  ptx::grid_dep_wait();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];\n" : "=r"(tmem_base) : "r"(tmem_ptr_smem));
  tmem_base = __shfl_sync(0xffffffffu, tmem_base, 0);

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (warp-wide,
    // one elected lane issues; see conv_patch.cu for why the roles are not run under lane == 0)
    uint32_t stage = 0, phase = 0;
    // my slice of the activation box: offset along the split dimension, and where it lands
    const int sl = items.rn * p.a_slice_step;
    const int sl_w = p.a_slice_dim == 0 ? sl : 0, sl_h = p.a_slice_dim == 1 ? sl : 0;
    const int sl_t = p.a_slice_dim == 2 ? sl : 0, sl_b = p.a_slice_dim == 3 ? sl : 0;
    const uint32_t a_dst = items.rn * a_slice_bytes;
    const uint32_t b_dst = a_bytes + items.rm * b_slice_bytes;
    const int b_row = items.rm * (p.block_n / p.cm);
    for (int item = first_item; item < items.total; item += item_step) {
      int m, nt;
      items.tile(item, &m, &nt);
      const TileCoord tc = decode_tile(p, m, nt);
      for (int tap = 0; tap < p.num_taps; ++tap) {
        const void* amap = &p.a_map[p.tap_map[tap]];
        const int cw = tc.w0 + p.tap_dw[tap] + sl_w;
        const int ch = tc.h0 + p.tap_dh[tap] + sl_h;
        const int ct = tc.t0 + p.tap_dt[tap] + sl_t;
        const int cb = tc.b0 + sl_b;
        const int coff = p.tap_coff[tap];
        const int koff = p.tap_koff[tap];
        for (int kb = 0; kb < p.tap_nkb[tap]; ++kb) {
          // free in EVERY CTA my slices go to (their MMA warps all arrive here)
          ptx::mbar_wait(empty_bar + 8 * stage, phase ^ 1);
          const uint32_t fb = full_bar + 8 * stage;
          const uint32_t sa = smem_base + stage * stage_bytes;
          ptx::mbar_expect_tx_elect(fb, stage_bytes);
          if (p.cn > 1)
            ptx::tma_load_5d_mcast_elect(sa + a_dst, amap, fb, coff + kb * p.bk, cw, ch, ct, cb,
                                         a_mask);
          else
            ptx::tma_load_5d_elect(sa, amap, fb, coff + kb * p.bk, cw, ch, ct, cb);
          if (p.cm > 1)
            ptx::tma_load_2d_mcast_elect(sa + b_dst, &p.b_map, fb, koff + kb * p.bk, tc.n0 + b_row,
                                         b_mask);
          else
            ptx::tma_load_2d_elect(sa + a_bytes, &p.b_map, fb, koff + kb * p.bk, tc.n0);
          if (++stage == (uint32_t)p.num_stages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (warp-wide)
    const uint32_t swz = p.bk * 2;  // bytes per operand row == swizzle span
    uint32_t stage = 0, phase = 0;
    int local = 0;
    for (int item = first_item; item < items.total; item += item_step, ++local) {
      // the issuing thread only needs the width of the channel tile, not the full tile decode
      const int nt = (item % items.n_groups) * items.cn + items.rn;
      const int ncols = min(p.block_n, p.cout_pad - nt * p.block_n);
      const bool live = ncols > 0;  // padding CTA of a cluster: keeps the barriers moving only
      const uint32_t acc = local & 1;
      const uint32_t acc_phase = (local >> 1) & 1;
      const uint32_t idesc = ptx::make_idesc_bf16_m128((uint32_t)(live ? ncols : 16));
      const uint32_t tmem_d = tmem_base + acc * kMaxBlockN;
      ptx::mbar_wait(tmem_empty_bar + 8 * acc, acc_phase ^ 1);
      ptx::tc_fence_after();
      uint32_t accumulate = 0;
      for (int tap = 0; tap < p.num_taps; ++tap) {
        for (int kb = 0; kb < p.tap_nkb[tap]; ++kb) {
          ptx::mbar_wait(full_bar + 8 * stage, phase);
          ptx::tc_fence_after();
          const uint32_t sa = smem_base + stage * stage_bytes;
          const uint64_t adesc = ptx::make_kmajor_desc(sa, swz);
          const uint64_t bdesc = ptx::make_kmajor_desc(sa + a_bytes, swz);
          const int ksteps = min(p.bk, p.tap_cin[tap] - kb * p.bk) >> 4;
          // advancing 16 bf16 = 32 bytes along K inside the swizzle span: +2 in the address field
          if (live) {
            ptx::umma_bf16_elect(tmem_d, adesc, bdesc, idesc, accumulate);
            if (ksteps > 1) ptx::umma_bf16_elect(tmem_d, adesc + 2, bdesc + 2, idesc, 1);
            if (ksteps > 2) ptx::umma_bf16_elect(tmem_d, adesc + 4, bdesc + 4, idesc, 1);
            if (ksteps > 3) ptx::umma_bf16_elect(tmem_d, adesc + 6, bdesc + 6, idesc, 1);
          }
          accumulate = 1;
          // the stage may be refilled once its readers are done: tell every CTA that writes into it
          if (csize > 1)
            ptx::umma_commit_mcast_elect(empty_bar + 8 * stage, peers_mask);
          else
            ptx::umma_commit_elect(empty_bar + 8 * stage);
          if (++stage == (uint32_t)p.num_stages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
      ptx::umma_commit_elect(tmem_full_bar + 8 * acc);
    }
  } else {
    // ------------------------------------------------------------------ epilogue (256 threads)
    // Two warps share each TMEM lane quarter and take alternate 16-column chunks.  Every thread
    // owns one output position (= TMEM lane) and writes its 16 channels as one 32-byte store; no
    // staging tile, so all shared memory goes to the operand ring.
    const int ew = warp - 2;          // 0..7
    const int q = warp & 3;           // TMEM lane quarter this warp may read
    const int half = ew >> 2;         // which chunks this warp converts
    const int row = q * 32 + lane;    // output position inside the tile
    const int et = threadIdx.x - 64;  // 0..255
    const int lbb_shift = p.lbw + p.lbh + p.lbt;
    const int iw = row & ((1 << p.lbw) - 1);
    const int ih = (row >> p.lbw) & ((1 << p.lbh) - 1);
    const int it = (row >> (p.lbw + p.lbh)) & ((1 << p.lbt) - 1);
    const int ib = row >> lbb_shift;
    int local = 0;
    int table_n0 = -1;
#ifdef RNB_PATCH_COUNTERS
    unsigned long long epi_cnt[6] = {};  // (only the patch kernel reports them)
#endif
    // this thread's slice of a work item: where its output row lives and whether it exists
    struct Slice {
      size_t off;
      int n0, ncols;
      bool valid;
    };
    auto slice_of = [&](int item) {
      int m, nt;
      items.tile(item, &m, &nt);
      const TileCoord tc = decode_tile(p, m, nt);
      const int w = tc.w0 + iw, h = tc.h0 + ih, t = tc.t0 + it, b = tc.b0 + ib;
      Slice s;
      s.valid = (w < p.Wo) && (h < p.Ho) && (t < p.To) && (b < p.B);
      s.off = ((((size_t)b * p.To + t) * p.Ho + h) * p.Wo + w) * p.cout_pad + tc.n0;
      s.n0 = tc.n0;
      s.ncols = tc.ncols > 0 ? tc.ncols : 0;
      return s;
    };
    // One work item.  The residual rows of the item after the next are prefetched into L2 before
    // the wait for this item's accumulator, two tile periods ahead of their use (one period is
    // about one loaded DRAM round trip, which is not enough).
    auto one_item = [&](int item, const Slice& cur, const Slice& nxt) {
      const uint32_t acc = local & 1;
      const uint32_t acc_phase = (local >> 1) & 1;
      Slice nxt2 = nxt;
      if (item + 2 * item_step < items.total) {
        nxt2 = slice_of(item + 2 * item_step);
        if (kRes) prefetch_residual_l2(p.residual + nxt2.off, nxt2.ncols, nxt2.valid);
      }
      if (cur.n0 != table_n0) {
        // per-channel shift of this channel tile -> shared memory.  The barrier before the write
        // keeps it behind the previous tile's readers, the one after publishes it.
        ptx::named_bar_sync(1, kEpiThreads);
        if (et < cur.ncols) {
          const float v = __ldg(p.shift + cur.n0 + et);
          asm volatile("st.shared.f32 [%0], %1;\n" ::"r"(shift_tab + et * 4), "f"(v) : "memory");
        }
        ptx::named_bar_sync(1, kEpiThreads);
        table_n0 = cur.n0;
      }
      const uint32_t taddr_row = tmem_base + ((uint32_t)(q * 32) << 16) + acc * kMaxBlockN;
      epilogue_tile<kRes, kRelu>(tmem_full_bar + 8 * acc, acc_phase, taddr_row, shift_tab,
                                 cur.ncols >> 4, half, cur.valid, p.y + cur.off,
                                 kRes ? p.residual + cur.off : nullptr
#ifdef RNB_PATCH_COUNTERS
                                 ,
                                 epi_cnt
#endif
      );
      // every TMEM read of this accumulator buffer has completed: hand it back to the MMA warp,
      // one arrival per warp
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(tmem_empty_bar + 8 * acc);
      ++local;
      return nxt2;
    };
    Slice cur = {}, nxt = {};
    if (first_item < items.total) {
      cur = slice_of(first_item);
      if (kRes) prefetch_residual_l2(p.residual + cur.off, cur.ncols, cur.valid);
    }
    if (first_item + item_step < items.total) {
      nxt = slice_of(first_item + item_step);
      if (kRes) prefetch_residual_l2(p.residual + nxt.off, nxt.ncols, nxt.valid);
    }
    for (int item = first_item; item < items.total; item += item_step) {
      const Slice nxt2 = one_item(item, cur, nxt);
      cur = nxt;
      nxt = nxt2;
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  // a peer may still be multicasting its last arrivals into my barriers
  if (csize > 1) ptx::cluster_sync_all();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, kTmemCols);
  }
}

namespace {
template <bool kRes, bool kRelu>
cudaError_t launch_variant(const ConvTcParams& p, int num_sms, int smem, cudaStream_t stream) {
  static std::atomic<bool> configured[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    e = cudaFuncSetAttribute(conv_tc_kernel<kRes, kRelu>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit);
    if (e != cudaSuccess) return e;
    configured[dev] = true;
  }
  const int csize = p.cm * p.cn;
  const int items = ((p.m_tiles + p.cm - 1) / p.cm) * ((p.n_tiles + p.cn - 1) / p.cn);
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(kConvThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[3];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int max_clusters = num_sms / csize;
  if (csize > 1) {
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = csize;
    attr[1].val.clusterDim.y = 1;
    attr[1].val.clusterDim.z = 1;
    cfg.numAttrs = 2;
    // how many clusters of this shape the device can hold at once (GPC boundaries strand SMs)
    static std::atomic<int> cached[64][5] = {};
    if (dev >= 0 && dev < 64 && cached[dev][csize] == 0) {
      cfg.gridDim = dim3(csize * (num_sms / csize));
      int n = 0;
      e = cudaOccupancyMaxActiveClusters(&n, conv_tc_kernel<kRes, kRelu>, &cfg);
      if (e != cudaSuccess) return e;
      cached[dev][csize] = n > 0 ? n : 1;
    }
    if (dev >= 0 && dev < 64 && cached[dev][csize] < max_clusters) max_clusters = cached[dev][csize];
  }
  const int clusters = items < max_clusters ? items : max_clusters;
  cfg.gridDim = dim3(clusters * csize);
  cfg.numAttrs = add_l2_output_window(attr, cfg.numAttrs, p.y,
                                      (size_t)p.B * p.To * p.Ho * p.Wo * p.cout_pad * 2);
  return cudaLaunchKernelEx(&cfg, conv_tc_kernel<kRes, kRelu>, p);
}
}  // namespace

namespace {
struct L2PersistState {
  bool init = false;
  size_t max_carve = 0;  // what the device (and RNB_B200_L2_PERSIST_MB) allow
  size_t carve = 0;      // current cudaLimitPersistingL2CacheSize set by this library (0 = off)
  int max_window = 0;
  size_t prev_limit = 0;  // the device's limit before this library first changed it
  bool touched = false;
  int users = 0;          // finalized model handles on this device
};
L2PersistState g_l2[64];

L2PersistState* l2_state() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  L2PersistState& st = g_l2[dev];
  if (!st.init) {
    st.init = true;
    // default: as much as the device allows; RNB_B200_L2_PERSIST_MB caps it, 0 = feature off
    const char* e = getenv("RNB_B200_L2_PERSIST_MB");
    const double mb = e ? atof(e) : 1e6;
    int max_bytes = 0;
    cudaDeviceGetAttribute(&max_bytes, cudaDevAttrMaxPersistingL2CacheSize, dev);
    cudaDeviceGetAttribute(&st.max_window, cudaDevAttrMaxAccessPolicyWindowSize, dev);
    size_t want = mb > 0 ? (size_t)(mb * 1e6) : 0;
    if (want > (size_t)max_bytes) want = (size_t)max_bytes;
    st.max_carve = want;
    (void)cudaGetLastError();
  }
  return &st;
}
}  // namespace

void l2_persist_configure(size_t largest_tensor_bytes) {
  L2PersistState* st = l2_state();
  if (!st || !st->max_carve) return;
  // The carve-out is taken away from ordinary L2 capacity.  It pays while the largest tensor of a
  // forward mostly fits (R18: +1.3 % at batch 16) and costs once the tensors outgrow it (-7 % at
  // batch 32, -19 % at batch 64 with the carve-out left on): measured crossover ~1.5x.
  const size_t want = (double)largest_tensor_bytes <= 1.5 * (double)st->max_carve ? st->max_carve : 0;
  if (want == st->carve) return;
  if (!st->touched) {
    // remembered so that the last handle's destruction can put the device-wide limit back
    if (cudaDeviceGetLimit(&st->prev_limit, cudaLimitPersistingL2CacheSize) != cudaSuccess)
      st->prev_limit = 0;
    st->touched = true;
  }
  if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want) == cudaSuccess) st->carve = want;
  (void)cudaGetLastError();
}

void l2_persist_acquire() {
  if (L2PersistState* st = l2_state()) ++st->users;
}

void l2_persist_release() {
  L2PersistState* st = l2_state();
  if (!st || st->users <= 0) return;
  if (--st->users == 0 && st->touched) {
    (void)cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, st->prev_limit);
    (void)cudaGetLastError();
    st->carve = 0;
    st->touched = false;
  }
}

int add_l2_output_window(cudaLaunchAttribute* attr, int n, const void* ptr, size_t bytes) {
  L2PersistState* st = l2_state();
  if (!st || !st->carve || !ptr || !bytes) return n;
  size_t win = bytes;
  if (st->max_window > 0 && win > (size_t)st->max_window) win = (size_t)st->max_window;
  attr[n].id = cudaLaunchAttributeAccessPolicyWindow;
  attr[n].val.accessPolicyWindow.base_ptr = const_cast<void*>(ptr);
  attr[n].val.accessPolicyWindow.num_bytes = win;
  const double ratio = (double)st->carve * 0.9 / (double)win;
  attr[n].val.accessPolicyWindow.hitRatio = ratio > 1.0 ? 1.0f : (float)ratio;
  attr[n].val.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
  attr[n].val.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
  return n + 1;
}

cudaError_t launch_conv_tc(const ConvTcParams& p, int num_sms, cudaStream_t stream) {
  const int smem = conv_tc_smem_bytes(p.bk, p.block_n, p.num_stages);
  if (p.cm < 1 || p.cn < 1 || p.cm * p.cn > 4) return cudaErrorInvalidValue;
  const bool res = p.residual != nullptr, relu = p.relu != 0;
  if (res) return relu ? launch_variant<true, true>(p, num_sms, smem, stream)
                       : launch_variant<true, false>(p, num_sms, smem, stream);
  return relu ? launch_variant<false, true>(p, num_sms, smem, stream)
              : launch_variant<false, false>(p, num_sms, smem, stream);
}

}  // namespace rnb
