// CTA-pair variant of the tap kernel (conv_tc.cu): tcgen05.mma.cta_group::2, M = 256.
//
// The tap kernel is bound by the bytes each SM has to pull in per tile (activation rows of every
// tap + the weight tile, against ~68 B/clk of inbound TMA per SM; profiles/r01_mma_issue_notes.md).
// Here two CTAs of a cluster (two SMs) work on two neighbouring position tiles and the SAME channel
// tile: each CTA loads its own 128 activation rows but only HALF of the weight tile, and the
// leader's tensor-core instruction reads both halves (its own shared memory and the peer's).  That
// halves the weight bytes per SM -- the larger operand in the conv3-conv5 layers.
//
//   warp 0 (both CTAs)  TMA producer: own activation box + own half of the weight tile per K block,
//                       completing on the LEADER's "full" barrier
//   warp 1 (leader)     issues the M=256 MMAs; its commits release the stage in both CTAs and
//                       publish the accumulator in both CTAs (multicast arrive)
//   warps 2-9 (both)    epilogue of the CTA's own 128 accumulator rows (its own TMEM), shared with
//                       the other kernels (epilogue.cuh); one arrival per CTA hands the accumulator
//                       buffer back to the leader
#include <atomic>
#include <stdio.h>

#include "conv_tc.cuh"
#include "epilogue.cuh"
#include "ptx.cuh"

namespace rnb {

#ifdef RNB_PATCH_COUNTERS
// development aid (make test_conv_cnt): cycles CTA 0's producer and MMA warps spend waiting
__device__ unsigned long long g_pair_cnt[8];
void pair_counters_read_reset(unsigned long long out[8]) {
  cudaMemcpyFromSymbol(out, g_pair_cnt, sizeof(g_pair_cnt));
  unsigned long long z[8] = {};
  cudaMemcpyToSymbol(g_pair_cnt, z, sizeof(z));
}
#define PCNT(i, stmt)                                              \
  {                                                                \
    const long long pc_t0 = clock64();                             \
    stmt;                                                          \
    pc_acc[i] += (unsigned long long)(clock64() - pc_t0);          \
  }
#else
#define PCNT(i, stmt) stmt
#endif

namespace {

struct PairTile {
  int w0, h0, t0, b0, n0, ncols;
};

// item = (pair of position tiles, channel tile), channel tile fastest; rank picks the position tile
// (the second of the last pair may lie past the last tile: its batch origin is then outside the
// tensor, its loads are zero-filled and its stores masked)
__device__ __forceinline__ PairTile pair_decode(const ConvTcParams& p, int item, int rank) {
  PairTile c;
  const int nt = item % p.n_tiles;
  int m = (item / p.n_tiles) * 2 + rank;
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

}  // namespace

template <bool kRes, bool kRelu>
__global__ void __launch_bounds__(kConvThreads, 1)
conv_pair_kernel(const __grid_constant__ ConvTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const uint32_t rank = blockIdx.x & 1u;  // (2,1,1) cluster over a 1-D grid; provably uniform

  // temporal-patch mode: tp_rows activation rows and three weight tiles (one per tap) per stage
  const uint32_t a_bytes = p.tpatch ? (uint32_t)p.tp_rows * 128u : kTileM * p.bk * 2;
  const uint32_t b_tile = (p.block_n / 2) * p.bk * 2;  // my half of one weight tile
  const uint32_t b_bytes = p.tpatch ? 3u * b_tile : b_tile;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  const uint32_t bars = smem_base + p.num_stages * stage_bytes;
  const uint32_t full_bar = bars;
  const uint32_t empty_bar = bars + 8 * p.num_stages;
  const uint32_t tmem_full_bar = bars + 16 * p.num_stages;
  const uint32_t tmem_empty_bar = tmem_full_bar + 16;
  const uint32_t tmem_ptr_smem = tmem_empty_bar + 16;
  const uint32_t shift_tab = bars + 256;
  // barriers the MMA warp waits on live in the leader
  const uint32_t full_ld = ptx::map_to_cta(full_bar, 0);
  const uint32_t tmem_empty_ld = ptx::map_to_cta(tmem_empty_bar, 0);

  const int total_items = ((p.m_tiles + 1) / 2) * p.n_tiles;
  const int first_item = blockIdx.x >> 1;
  const int item_step = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&p.a_map[0]);
    ptx::prefetch_tmap(&p.b_map);
    for (int s = 0; s < p.num_stages; ++s) {
      ptx::mbar_init(full_bar + 8 * s, 1);
      ptx::mbar_init(empty_bar + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(tmem_full_bar + 8 * s, 1);
      ptx::mbar_init(tmem_empty_bar + 8 * s, 2);  // one arrival per CTA
    }
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc_pair(tmem_ptr_smem, kTmemCols);
    ptx::tmem_relinquish_pair();
  }
  ptx::tc_fence_before();
  __syncthreads();
  // the peer signals the leader's barriers and the leader's commits arrive on the peer's: nobody
  // starts before both CTAs have initialised theirs
  ptx::cluster_sync_all();
  ptx::tc_fence_after();
  ptx::grid_dep_launch_dependents();
  ptx::grid_dep_wait();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];\n" : "=r"(tmem_base) : "r"(tmem_ptr_smem));
  tmem_base = __shfl_sync(0xffffffffu, tmem_base, 0);

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (warp-wide)
    uint32_t stage = 0, phase = 0;
#ifdef RNB_PATCH_COUNTERS
    unsigned long long pc_acc[8] = {};
#endif
    for (int item = first_item; item < total_items; item += item_step) {
      const PairTile tc = pair_decode(p, item, rank);
      // my rows of the channel tile: the MMA takes columns [0, N/2) from the leader's shared
      // memory and [N/2, N) from the peer's
      const int brow = tc.n0 + rank * (tc.ncols / 2);
      if (p.tpatch) {
        for (int kb = 0; kb < p.kb_per_tap; ++kb) {
          PCNT(0, ptx::mbar_wait(empty_bar + 8 * stage, phase ^ 1));
          const uint32_t fb = full_ld + 8 * stage;
          const uint32_t sa = smem_base + stage * stage_bytes;
          if (rank == 0) ptx::mbar_expect_tx_elect(full_bar + 8 * stage, 2 * stage_bytes);
          // frames t0 - 1 .. t0 + tile frames: the halo lies outside the tensor at the clip's ends
          // and is zero-filled, which is the convolution's temporal padding
          ptx::tma_load_5d_pair_elect(sa, &p.a_map[0], fb, kb * p.bk, tc.w0, 0, tc.t0 - 1, tc.b0);
          for (int tap = 0; tap < 3; ++tap)
            ptx::tma_load_2d_pair_elect(sa + a_bytes + tap * b_tile, &p.b_map, fb,
                                        tap * p.cin_pad + kb * p.bk, brow);
          if (++stage == (uint32_t)p.num_stages) {
            stage = 0;
            phase ^= 1;
          }
        }
        continue;
      }
      for (int tap = 0; tap < p.num_taps; ++tap) {
        const void* amap = &p.a_map[p.tap_map[tap]];
        const int cw = tc.w0 + p.tap_dw[tap], ch = tc.h0 + p.tap_dh[tap], ct = tc.t0 + p.tap_dt[tap];
        const int coff = p.tap_coff[tap], koff = p.tap_koff[tap];
        for (int kb = 0; kb < p.tap_nkb[tap]; ++kb) {
          PCNT(0, ptx::mbar_wait(empty_bar + 8 * stage, phase ^ 1));
          const uint32_t fb = full_ld + 8 * stage;
          const uint32_t sa = smem_base + stage * stage_bytes;
          // the leader announces the bytes of both CTAs
          if (rank == 0) ptx::mbar_expect_tx_elect(full_bar + 8 * stage, 2 * stage_bytes);
          ptx::tma_load_5d_pair_elect(sa, amap, fb, coff + kb * p.bk, cw, ch, ct, tc.b0);
          ptx::tma_load_2d_pair_elect(sa + a_bytes, &p.b_map, fb, koff + kb * p.bk, brow);
          if (++stage == (uint32_t)p.num_stages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
#ifdef RNB_PATCH_COUNTERS
    if (blockIdx.x == 0 && lane == 0) g_pair_cnt[0] += pc_acc[0];
#endif
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (leader only)
    if (rank == 0) {
      const uint32_t swz = p.bk * 2;
      uint32_t stage = 0, phase = 0;
      int local = 0;
#ifdef RNB_PATCH_COUNTERS
      unsigned long long pc_acc[8] = {};
      const long long pc_loop0 = clock64();
      unsigned long long pc_g0;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(pc_g0));
#endif
      for (int item = first_item; item < total_items; item += item_step, ++local) {
        const int ncols = min(p.block_n, p.cout_pad - (item % p.n_tiles) * p.block_n);
        const uint32_t idesc = ptx::make_idesc_bf16_m256((uint32_t)ncols);
        const uint32_t acc = local & 1;
        const uint32_t acc_phase = (local >> 1) & 1;
        const uint32_t tmem_d = tmem_base + acc * kMaxBlockN;
        PCNT(1, ptx::mbar_wait(tmem_empty_bar + 8 * acc, acc_phase ^ 1));
        ptx::tc_fence_after();
        uint32_t accumulate = 0;
        if (p.tpatch) {
          for (int kb = 0; kb < p.kb_per_tap; ++kb) {
            PCNT(2, ptx::mbar_wait(full_bar + 8 * stage, phase));
            ptx::tc_fence_after();
            const uint32_t sa = smem_base + stage * stage_bytes;
            const int ksteps = min(p.bk, p.cin_pad - kb * p.bk) >> 4;
            for (int tap = 0; tap < 3; ++tap) {
              // tap dt = the same rows, tp_shift rows (one frame of the tile) further on
              const uint64_t adesc =
                  ptx::make_kmajor_desc(sa + (uint32_t)(tap * p.tp_shift) * 128u, swz);
              const uint64_t bdesc = ptx::make_kmajor_desc(sa + a_bytes + tap * b_tile, swz);
              ptx::umma_bf16_pair_elect(tmem_d, adesc, bdesc, idesc, accumulate);
              if (ksteps > 1) ptx::umma_bf16_pair_elect(tmem_d, adesc + 2, bdesc + 2, idesc, 1);
              if (ksteps > 2) ptx::umma_bf16_pair_elect(tmem_d, adesc + 4, bdesc + 4, idesc, 1);
              if (ksteps > 3) ptx::umma_bf16_pair_elect(tmem_d, adesc + 6, bdesc + 6, idesc, 1);
              accumulate = 1;
            }
            ptx::umma_commit_pair_elect(empty_bar + 8 * stage, 3);
            if (++stage == (uint32_t)p.num_stages) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
        for (int tap = 0; tap < (p.tpatch ? 0 : p.num_taps); ++tap) {
          for (int kb = 0; kb < p.tap_nkb[tap]; ++kb) {
            PCNT(2, ptx::mbar_wait(full_bar + 8 * stage, phase));
            ptx::tc_fence_after();
            const uint32_t sa = smem_base + stage * stage_bytes;
            const uint64_t adesc = ptx::make_kmajor_desc(sa, swz);
            const uint64_t bdesc = ptx::make_kmajor_desc(sa + a_bytes, swz);
            const int ksteps = min(p.bk, p.tap_cin[tap] - kb * p.bk) >> 4;
            ptx::umma_bf16_pair_elect(tmem_d, adesc, bdesc, idesc, accumulate);
            if (ksteps > 1) ptx::umma_bf16_pair_elect(tmem_d, adesc + 2, bdesc + 2, idesc, 1);
            if (ksteps > 2) ptx::umma_bf16_pair_elect(tmem_d, adesc + 4, bdesc + 4, idesc, 1);
            if (ksteps > 3) ptx::umma_bf16_pair_elect(tmem_d, adesc + 6, bdesc + 6, idesc, 1);
            accumulate = 1;
            ptx::umma_commit_pair_elect(empty_bar + 8 * stage, 3);  // both CTAs may refill
            if (++stage == (uint32_t)p.num_stages) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
        ptx::umma_commit_pair_elect(tmem_full_bar + 8 * acc, 3);  // both epilogues may start
      }
#ifdef RNB_PATCH_COUNTERS
      if (blockIdx.x == 0 && lane == 0) {
        unsigned long long pc_g1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(pc_g1));
        g_pair_cnt[1] += pc_acc[1];
        g_pair_cnt[2] += pc_acc[2];
        g_pair_cnt[3] += (unsigned long long)local;
        g_pair_cnt[6] += (unsigned long long)(clock64() - pc_loop0);
        g_pair_cnt[7] += pc_g1 - pc_g0;
      }
#endif
    }
  } else {
    // ------------------------------------------------------------------ epilogue (256 threads)
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
    int local = 0;
    int table_n0 = -1;
#ifdef RNB_PATCH_COUNTERS
    unsigned long long epi_cnt[6] = {};
#endif
    struct Slice {
      size_t off;
      int n0, ncols;
      bool valid;
    };
    auto slice_of = [&](int item) {
      const PairTile tc = pair_decode(p, item, rank);
      const int w = tc.w0 + iw, h = tc.h0 + ih, t = tc.t0 + it, b = tc.b0 + ib;
      Slice s;
      s.valid = (w < p.Wo) && (h < p.Ho) && (t < p.To) && (b < p.B);
      s.off = ((((size_t)b * p.To + t) * p.Ho + h) * p.Wo + w) * p.cout_pad + tc.n0;
      s.n0 = tc.n0;
      s.ncols = tc.ncols;
      return s;
    };
    Slice cur = {};
    if (first_item < total_items) cur = slice_of(first_item);
    for (int item = first_item; item < total_items; item += item_step, ++local) {
      const uint32_t acc = local & 1;
      const uint32_t acc_phase = (local >> 1) & 1;
      Slice nxt = cur;
      if (item + item_step < total_items) {
        nxt = slice_of(item + item_step);
        if (kRes) prefetch_residual_l2(p.residual + nxt.off, nxt.ncols, nxt.valid);
      }
      if (cur.n0 != table_n0) {
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
      // all of this CTA's reads of the accumulator buffer are done: one arrival per CTA on the
      // leader's barrier
      ptx::tc_fence_before();
      ptx::named_bar_sync(2, kEpiThreads);
      if (et == 0) ptx::mbar_arrive_cluster(tmem_empty_ld + 8 * acc);
      cur = nxt;
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  // the leader's last commits may still be on their way to the peer's barriers
  ptx::cluster_sync_all();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc_pair(tmem_base, kTmemCols);
  }
}

namespace {
template <bool kRes, bool kRelu>
cudaError_t launch_pair_variant(const ConvTcParams& p, int num_sms, int smem, cudaStream_t stream) {
  static std::atomic<bool> configured[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    e = cudaFuncSetAttribute(conv_pair_kernel<kRes, kRelu>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit);
    if (e != cudaSuccess) return e;
    configured[dev] = true;
  }
  const int items = ((p.m_tiles + 1) / 2) * p.n_tiles;
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(kConvThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[3];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  attr[1].id = cudaLaunchAttributeClusterDimension;
  attr[1].val.clusterDim.x = 2;
  attr[1].val.clusterDim.y = 1;
  attr[1].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  // a persistent grid must be fully resident; not every SM has a partner for a 2-CTA cluster
  static std::atomic<int> cached[64] = {};
  int max_pairs = num_sms / 2;
  if (dev >= 0 && dev < 64) {
    if (cached[dev] == 0) {
      cfg.gridDim = dim3(2 * (num_sms / 2));
      int n = 0;
      e = cudaOccupancyMaxActiveClusters(&n, conv_pair_kernel<kRes, kRelu>, &cfg);
      if (e != cudaSuccess) return e;
      cached[dev] = n > 0 ? n : 1;
    }
    if (cached[dev] < max_pairs) max_pairs = cached[dev];
  }
  const int pairs = items < max_pairs ? items : max_pairs;
  cfg.gridDim = dim3(2 * pairs);
  cfg.numAttrs = add_l2_output_window(attr, cfg.numAttrs, p.y,
                                      (size_t)p.B * p.To * p.Ho * p.Wo * p.cout_pad * 2);
  return cudaLaunchKernelEx(&cfg, conv_pair_kernel<kRes, kRelu>, p);
}
}  // namespace

cudaError_t launch_conv_pair(const ConvTcParams& p, int num_sms, cudaStream_t stream) {
  if (!p.pair || p.cm != 1 || p.cn != 1 || p.ks > 1 || (p.block_n & 15)) return cudaErrorInvalidValue;
  const int smem =
      p.tpatch ? 1024 + p.num_stages * conv_tpatch_stage_bytes(p.tp_rows, p.block_n) + 256 + kMaxBlockN * 4
               : conv_pair_smem_bytes(p.bk, p.block_n, p.num_stages);
  if (smem > kSmemLimit) return cudaErrorInvalidValue;
  const bool res = p.residual != nullptr, relu = p.relu != 0;
  if (res && relu) return launch_pair_variant<true, true>(p, num_sms, smem, stream);
  if (res) return launch_pair_variant<true, false>(p, num_sms, smem, stream);
  if (relu) return launch_pair_variant<false, true>(p, num_sms, smem, stream);
  return launch_pair_variant<false, false>(p, num_sms, smem, stream);
}

}  // namespace rnb
