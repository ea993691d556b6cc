// Patch convolution (see conv_patch.cuh): one activation patch per tile, filter taps as
// row-shifted UMMA descriptors, weights resident in shared memory or streamed through a ring,
// epilogue straight from TMEM to global memory.
//
//   warp 0      TMA producer: (resident mode) all weight tiles once; per work item one patch
//               (nkb boxes, double buffered) and (stream mode) num_taps x nkb weight tiles
//   warp 1      TMEM owner and MMA issuer.  The whole warp runs the issue loop converged and one
//               elected lane issues each tcgen05.mma: with a shuffle-broadcast warp index / TMEM
//               address and all-lane mbarrier waits, ptxas keeps every descriptor in uniform
//               registers (back-to-back UTCHMMA).  Running the role inside `if (lane == 0)`, or
//               waiting with one lane + __syncwarp, costs ~120-146 cycles of R2UR traffic per MMA
//               against 72 cycles of tensor time at N=144 (profiles/r01_mma_issue_notes.md).
//   warps 2-9   epilogue: tcgen05.ld -> +shift (+residual) -> ReLU -> bf16 -> 32-byte global stores
#include <atomic>
#include <stdio.h>

#include "conv_patch.cuh"
#include "epilogue.cuh"
#include "ptx.cuh"

namespace rnb {

namespace {

struct PatchItem {
  int w0, h0, t0, b;  // position tile
  int n0, ncols;      // channel tile
};

// work items: channel tile fastest, then the position tiles in (w, h, t, b) order.  In pair mode an
// item covers two consecutive position tiles, one per CTA of the pair (`rank`); the second may lie
// past the last tile, its batch index is then out of range and everything it touches is masked.
__device__ __forceinline__ PatchItem decode_patch_item(const ConvPatchParams& p, int item, int pair,
                                                       int rank) {
  PatchItem c;
  const int nt = item % p.n_tiles;
  int tile = (item / p.n_tiles) * pair + rank;
  const int iw = tile % p.tiles_w;
  tile /= p.tiles_w;
  const int ih = tile % p.tiles_h;
  tile /= p.tiles_h;
  const int it = tile % p.tiles_t;
  c.b = tile / p.tiles_t;
  if (p.reverse && c.b < p.tiles_b) c.b = p.tiles_b - 1 - c.b;
  c.w0 = iw * p.tw;
  c.h0 = ih * p.th;
  c.t0 = it * p.tt;
  c.n0 = nt * p.block_n;
  c.ncols = min(p.block_n, p.n - c.n0);
  return c;
}

}  // namespace

// kPair: two CTAs of a cluster issue ONE tcgen05.mma.cta_group::2 (M = 256): each CTA holds its own
// position tile (patch, accumulator, epilogue) and HALF of every weight tile.  That halves the
// weight-side shared-memory reads and fills, which is what caps the single-CTA kernel (UMMA operand
// reads 64*(M+N)/N B/clk plus TMA fills against a 128 B/clk port).  The leader (rank 0) issues all
// MMAs; both CTAs' TMA loads signal the leader's "full" barriers, the leader's commits release the
// buffers and publish the accumulators in both CTAs.
#ifdef RNB_PATCH_COUNTERS
// development aid (make test_conv_cnt): cycles CTA 0's warps spend waiting, summed over its tiles
__device__ unsigned long long g_patch_cnt[8];
__device__ unsigned long long g_patch_epi[8];
__device__ unsigned long long g_patch_cta[4][3][256];
__device__ unsigned int g_patch_smid[4][256];  // per CTA: globaltimer at entry, MMA loop start/end
__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// (accumulated in registers, written once when the role ends: a global read-modify-write per
// wait would itself stall the warp being measured)
#define CNT_DECL() unsigned long long cnt_acc[6] = {}
#define CNT_BEGIN() const long long cnt_t0 = clock64()
#define CNT_END(i) cnt_acc[i] += (unsigned long long)(clock64() - cnt_t0)
#define CNT_FLUSH(i) \
  if (blockIdx.x == 0 && lane == 0) g_patch_cnt[i] += cnt_acc[i]
#else
#define CNT_DECL()
#define CNT_BEGIN()
#define CNT_END(i)
#define CNT_FLUSH(i)
#endif

// One PHASE of a patch-kernel CTA: a contiguous range of work items of ONE convolution, taken
// round-robin by the `item_step` CTAs (or CTA pairs) that run the phase.  The plain kernel is a
// single phase over all items.  The co-scheduled kernel (conv_patch_dual_kernel below) runs the
// spatial convolution of a SpatioTemporalConv on some CTAs and its temporal convolution on the
// others in the same launch, two or three phases per CTA.
struct PatchPhase {
  int first_item, item_step;  // this CTA's (pair's) first item inside the range, and the stride
  int item_lo, item_hi;       // the range of items of the phase
  int bar_slot;               // which of the kPatchBarSlots barrier areas the phase uses
  int load_weights;           // 0: the resident weights are still those of the previous phase
  // Producer side of the per-clip dependency: sig[b] is bumped once per finished item of clip b.
  unsigned long long* sig;
  // Consumer side: an item of clip b may load its patch once dep[b] >= dep_target.
  const unsigned long long* dep;
  unsigned long long dep_target;
  // Dynamic phases (kDyn): instead of the fixed round-robin, the CTA's producer warp claims
  // `chunk` consecutive items at a time from the grid-wide counter *claim (items item_lo + claim
  // value ...) and hands each claim to the other warps through a two-entry mailbox, so CTAs that
  // join late or run slower simply take fewer items.
  unsigned long long* claim;
  int chunk;
};

constexpr int kPatchBarSlots = 3;
constexpr uint32_t kPatchBarSlotBytes = 320;
// control block at the start of shared memory: barrier areas, shift table, TMEM address
constexpr uint32_t kPatchCtlBytes = 2048;
static_assert(kPatchBarSlots * kPatchBarSlotBytes + kMaxBlockN * 4 + 16 <= kPatchCtlBytes,
              "control block");

template <bool kRes, bool kRelu, bool kStream, bool kPair, bool kDyn = false>
__device__ __forceinline__ void conv_patch_phase(const ConvPatchParams& p, const PatchPhase ph,
                                                 uint32_t smem_base, uint32_t tmem_base, int warp,
                                                 int lane, uint32_t rank) {
  constexpr int kNumCta = kPair ? 2 : 1;
  // one (tap, channel block) weight tile, or this CTA's half of it
  const uint32_t wtile_bytes = (p.block_n / kNumCta) * 128;
  // tail block (resident single-CTA mode only): rows of tail_row bytes, swizzled at that width
  const uint32_t tail_row = (!kStream && !kPair) ? (uint32_t)p.tail_ch * 2 : 0u;
  const uint32_t wtail_bytes = p.block_n * tail_row;
  const uint32_t wtap_bytes = p.nkb * wtile_bytes + wtail_bytes;  // all weight tiles of one tap
  const uint32_t pblock_bytes = p.patch_rows * 128;     // one channel block of a patch
  const uint32_t ptail_off = p.nkb * pblock_bytes;      // the tail block follows the full ones
  const uint32_t patch_bytes = ptail_off + p.patch_rows * tail_row;
  const uint32_t bars = smem_base + (uint32_t)ph.bar_slot * kPatchBarSlotBytes;
  const uint32_t shift_tab = smem_base + kPatchBarSlots * kPatchBarSlotBytes;
  const uint32_t mailbox = shift_tab + kMaxBlockN * 4u + 8u;  // [2] claimed first items
  const uint32_t wbuf = smem_base + kPatchCtlBytes;
  const uint32_t pbuf =
      wbuf + (kStream ? (uint32_t)p.b_stages * wtile_bytes : (uint32_t)p.num_taps * wtap_bytes);
  // +1024: rows past the last patch that only garbage rows read; then two residual tiles
  const uint32_t rbuf = pbuf + p.p_bufs * patch_bytes + 1024;
  const bool res_smem = kRes && !kStream && !kPair && p.res_smem;
  const uint32_t rtile_bytes = res_smem ? (uint32_t)(p.n >> 6) * 16384u : 0u;
  const uint32_t w_full = bars;
  const uint32_t p_full = bars + 8;                     // [kPatchMaxBufs]
  const uint32_t p_empty = p_full + 8 * kPatchMaxBufs;  // [kPatchMaxBufs]
  const uint32_t tmem_full_bar = p_empty + 8 * kPatchMaxBufs;  // [2]
  const uint32_t tmem_empty_bar = tmem_full_bar + 16;          // [2]
  const uint32_t b_full = tmem_empty_bar + 16 + 8;            // [kPatchMaxBStages]
  const uint32_t b_empty = b_full + 8 * kPatchMaxBStages;     // [kPatchMaxBStages]
  const uint32_t mail_full = b_empty + 8 * kPatchMaxBStages;  // [2]
  const uint32_t mail_empty = mail_full + 16;                 // [2]
  static_assert(8 + 16 * kPatchMaxBufs + 32 + 8 + 16 * kPatchMaxBStages + 32 <= kPatchBarSlotBytes,
                "barrier area");
  // resident mode uses the stream barriers for its taps (<= 12 with res_smem) and the last four
  // for the residual tiles
  const uint32_t res_full = b_empty + 8 * (kPatchMaxBStages - 4);   // [2]
  const uint32_t res_empty = b_empty + 8 * (kPatchMaxBStages - 2);  // [2]

  const int total_items = ph.item_hi;
  const int first_item = ph.item_lo + ph.first_item;
  const int item_step = ph.item_step;
  // barriers the MMA warp waits on live in the leader; both CTAs' loads signal them
  const uint32_t p_full_ld = kPair ? ptx::map_to_cta(p_full, 0) : p_full;
  const uint32_t b_full_ld = kPair ? ptx::map_to_cta(b_full, 0) : b_full;
  const uint32_t tmem_empty_ld = kPair ? ptx::map_to_cta(tmem_empty_bar, 0) : tmem_empty_bar;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&p.a_map);
    ptx::prefetch_tmap(&p.b_map);
    if (tail_row) {
      ptx::prefetch_tmap(&p.a_map_tail);
      ptx::prefetch_tmap(&p.b_map_tail);
    }
    ptx::mbar_init(w_full, 1);
    for (int s = 0; s < kPatchMaxBufs; ++s) {
      ptx::mbar_init(p_full + 8 * s, 1);
      ptx::mbar_init(p_empty + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(tmem_full_bar + 8 * s, 1);
      // pair mode: one arrival per CTA (its epilogue warps meet on a named barrier first) --
      // hundreds of remote arrivals per tile on one mbarrier serialise
      ptx::mbar_init(tmem_empty_bar + 8 * s, kPair ? 2 : kEpiThreads / 32);
    }
    for (int s = 0; s < kPatchMaxBStages; ++s) {
      ptx::mbar_init(b_full + 8 * s, 1);
      // (the residual "empty" barriers take one arrival per epilogue warp)
      ptx::mbar_init(b_empty + 8 * s, (res_smem && s >= kPatchMaxBStages - 2) ? kEpiThreads / 32 : 1);
    }
    if (kDyn)
      for (int s = 0; s < 2; ++s) {
        ptx::mbar_init(mail_full + 8 * s, 1);
        ptx::mbar_init(mail_empty + 8 * s, 1 + kEpiThreads / 32);  // MMA warp + every epilogue warp
      }
    ptx::fence_mbar_init();
  }
  if (!kStream && threadIdx.x >= 64) {
    // one channel tile for the whole phase: its shift table is loaded once
    const int et = threadIdx.x - 64;
    if (et < p.n) {
      const float v = __ldg(p.shift + et);
      asm volatile("st.shared.f32 [%0], %1;\n" ::"r"(shift_tab + et * 4), "f"(v) : "memory");
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  // the peer signals the leader's barriers and the leader's commits arrive on the peer's: nobody
  // starts before both CTAs have initialised theirs
  if (kPair) ptx::cluster_sync_all();
  ptx::tc_fence_after();

  // Dynamic phases: the consumer side of the mailbox (MMA warp and epilogue warps).  Returns the
  // first item of the next claim; the shuffle keeps the value provably warp-uniform.
  uint32_t mk = 0;      // claims seen so far by this warp
  int dyn_left = 0;     // items left in the current claim
  auto take_claim = [&]() -> int {
    ptx::mbar_wait(mail_full + 8 * (mk & 1), (mk >> 1) & 1);
    int base;
    asm volatile("ld.shared.b32 %0, [%1];\n" : "=r"(base) : "r"(mailbox + 4 * (mk & 1)) : "memory");
    base = __shfl_sync(0xffffffffu, base, 0);
    __syncwarp();
    if (lane == 0) ptx::mbar_arrive(mail_empty + 8 * (mk & 1));
    ++mk;
    return base;
  };

  CNT_DECL();
  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (warp-wide)
    if (!kStream && ph.load_weights) {
      // the packed weights are constants of the model: fetch them while the previous kernel in
      // the stream is still draining, and only then wait for its activations.  In pair mode the
      // leader announces the bytes of BOTH halves; each CTA fetches its own rows.
      // One barrier per tap (the stream-mode weight barriers are free in this mode): the first
      // tile starts on tap 0 while the other taps are still arriving, instead of idling for the
      // ~2 us the whole weight set takes at the SM's inbound TMA rate.
      const int row0 = rank * (p.n / kNumCta);
      for (int tap = 0; tap < p.num_taps; ++tap) {
        const uint32_t wt = wbuf + tap * wtap_bytes;
        if (rank == 0) ptx::mbar_expect_tx_elect(b_full + 8 * tap, kNumCta * wtap_bytes);
        for (int kb = 0; kb < p.nkb; ++kb) {
          if (kPair)
            ptx::tma_load_2d_pair_elect(wt + kb * wtile_bytes, &p.b_map, b_full_ld + 8 * tap,
                                        tap * p.cin_pad + kb * 64, row0);
          else
            ptx::tma_load_2d_elect(wt + kb * wtile_bytes, &p.b_map, b_full + 8 * tap,
                                   tap * p.cin_pad + kb * 64, 0);
        }
        if (tail_row)
          ptx::tma_load_2d_elect(wt + p.nkb * wtile_bytes, &p.b_map_tail, b_full + 8 * tap,
                                 tap * p.cin_pad + p.nkb * 64, 0);
      }
    }
    ptx::grid_dep_wait();
    uint32_t pb = 0, pph = 0;
    uint32_t bs = 0, bphase = 0;
    uint32_t rlocal = 0;
    int ready_b = -1;  // clips known to be complete (co-scheduled consumer phases)
    // dynamic phases: claim the next `chunk` items and post the claim for the other warps
    auto claim_next = [&]() -> int {
      ptx::mbar_wait(mail_empty + 8 * (mk & 1), ((mk >> 1) & 1) ^ 1);
      int base = 0;
      if (lane == 0) {
        base = ph.item_lo + (int)atomicAdd(ph.claim, (unsigned long long)ph.chunk);
        asm volatile("st.shared.b32 [%0], %1;\n" ::"r"(mailbox + 4 * (mk & 1)), "r"(base) : "memory");
        ptx::mbar_arrive(mail_full + 8 * (mk & 1));
      }
      base = __shfl_sync(0xffffffffu, base, 0);
      ++mk;
      return base;
    };
    int item = first_item;
    if (kDyn) {
      item = claim_next();
      dyn_left = ph.chunk;
    }
    for (; item < total_items;) {
      const PatchItem tc = decode_patch_item(p, item, kNumCta, rank);
      {
        CNT_BEGIN();
        ptx::mbar_wait(p_empty + 8 * pb, pph ^ 1);
        CNT_END(0);
      }
      if (ph.dep != nullptr && tc.b > ready_b) {
        // The patch is cut from a tensor that other CTAs of THIS launch are still writing: wait
        // until every tile of clip b has been stored (acquire), then order the TMA engine's reads
        // (async proxy) behind what the acquire made visible.
        if (lane == 0) {
          const unsigned long long* f = ph.dep + tc.b;
          unsigned long long v;
          const long long t0 = clock64();
          do {
            asm volatile("ld.acquire.gpu.global.u64 %0, [%1];\n" : "=l"(v) : "l"(f) : "memory");
            if (v < ph.dep_target && clock64() - t0 > 4000000000LL) {
              printf("rnb_b200: co-scheduled convolution: clip %d never completed (block %d)\n",
                     tc.b, (int)blockIdx.x);
              __trap();
            }
          } while (v < ph.dep_target);
        }
        __syncwarp();
        asm volatile("fence.proxy.async.global;\n" ::: "memory");
        ready_b = tc.b;
      }
      if (rank == 0) ptx::mbar_expect_tx_elect(p_full + 8 * pb, kNumCta * patch_bytes);
      for (int kb = 0; kb < p.nkb; ++kb) {
        const uint32_t dst = pbuf + pb * patch_bytes + kb * pblock_bytes;
        if (kPair)
          ptx::tma_load_5d_pair_elect(dst, &p.a_map, p_full_ld + 8 * pb, kb * 64, tc.w0 + p.aw,
                                      tc.h0 + p.ah, tc.t0 + p.at, tc.b);
        else
          ptx::tma_load_5d_elect(dst, &p.a_map, p_full + 8 * pb, kb * 64, tc.w0 + p.aw,
                                 tc.h0 + p.ah, tc.t0 + p.at, tc.b);
      }
      if (tail_row)
        ptx::tma_load_5d_elect(pbuf + pb * patch_bytes + ptail_off, &p.a_map_tail, p_full + 8 * pb,
                               p.nkb * 64, tc.w0 + p.aw, tc.h0 + p.ah, tc.t0 + p.at, tc.b);
      if (res_smem) {
        // the residual rows of this item's output tile, into the tile buffer of its accumulator
        const uint32_t acc = rlocal & 1, rphase = (rlocal >> 1) & 1;
        ptx::mbar_wait(res_empty + 8 * acc, rphase ^ 1);
        ptx::mbar_expect_tx_elect(res_full + 8 * acc, rtile_bytes);
        for (int rb = 0; rb < (p.n >> 6); ++rb)
          ptx::tma_load_5d_elect(rbuf + acc * rtile_bytes + rb * 16384, &p.r_map,
                                 res_full + 8 * acc, rb * 64, tc.w0, tc.h0, tc.t0, tc.b);
        ++rlocal;
      }
      if (kStream) {
        // my rows of the channel tile: the MMA takes columns [0, N/2) from the leader's shared
        // memory and [N/2, N) from the peer's
        const int brow = tc.n0 + rank * (tc.ncols / kNumCta);
        for (int tap = 0; tap < p.num_taps; ++tap)
          for (int kb = 0; kb < p.nkb; ++kb) {
            ptx::mbar_wait(b_empty + 8 * bs, bphase ^ 1);
            if (rank == 0) ptx::mbar_expect_tx_elect(b_full + 8 * bs, kNumCta * wtile_bytes);
            if (kPair)
              ptx::tma_load_2d_pair_elect(wbuf + bs * wtile_bytes, &p.b_map, b_full_ld + 8 * bs,
                                          tap * p.cin_pad + kb * 64, brow);
            else
              ptx::tma_load_2d_elect(wbuf + bs * wtile_bytes, &p.b_map, b_full + 8 * bs,
                                     tap * p.cin_pad + kb * 64, brow);
            if (++bs == (uint32_t)p.b_stages) {
              bs = 0;
              bphase ^= 1;
            }
          }
      }
      if (++pb == (uint32_t)p.p_bufs) {
        pb = 0;
        pph ^= 1;
      }
      if (kDyn) {
        if (--dyn_left > 0) {
          ++item;
        } else {
          item = claim_next();
          dyn_left = ph.chunk;
        }
      } else {
        item += item_step;
      }
    }
    CNT_FLUSH(0);
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (warp-wide)
    // (reads only shared memory and TMEM: ordered behind the producer's grid_dep_wait by the
    // patch-full barrier).  In pair mode only the leader issues.
    if (!kPair || rank == 0) {
#ifdef RNB_PATCH_COUNTERS
      if (lane == 0 && blockIdx.x < 256) g_patch_cta[p.dbg_launch & 3][1][blockIdx.x] = gtimer();
      const long long loop_c0 = clock64();
      const unsigned long long loop_g0 = gtimer();
#endif
      int local = 0;
      uint32_t pb = 0, pph = 0;
      uint32_t bs = 0, bphase = 0;
      // (Waiting for the NEXT tile's accumulator and patch half-way through the current tile, to
      // take the ~450-cycle barrier round trip off the tile boundary, was measured slower: the
      // previous tile's epilogue is not done by then and the MMA queue drains while the issuing
      // thread blocks.  profiles/r01_mma_issue_notes.md)
      int item = first_item;
      if (kDyn) {
        item = take_claim();
        dyn_left = ph.chunk;
      }
      for (; item < total_items; ++local) {
        // the issuing thread only needs the width of the channel tile, not the full tile decode
        // (its integer divisions would sit on the critical path of every tile boundary)
        const int ncols = kStream ? min(p.block_n, p.n - (item % p.n_tiles) * p.block_n) : p.n;
        const uint32_t idesc = kPair ? ptx::make_idesc_bf16_m256((uint32_t)ncols)
                                     : ptx::make_idesc_bf16_m128((uint32_t)ncols);
        const uint32_t acc = local & 1;
        const uint32_t acc_phase = (local >> 1) & 1;
        const uint32_t tmem_d = tmem_base + acc * kMaxBlockN;
        {
          CNT_BEGIN();
          ptx::mbar_wait(tmem_empty_bar + 8 * acc, acc_phase ^ 1);
          CNT_END(1);
        }
        {
          CNT_BEGIN();
          ptx::mbar_wait(p_full + 8 * pb, pph);
          CNT_END(2);
        }
        ptx::tc_fence_after();
        CNT_BEGIN();
        uint32_t accumulate = 0;
        uint32_t wt = wbuf;
        for (int tap = 0; tap < p.num_taps; ++tap) {
          const uint32_t arow = pbuf + pb * patch_bytes + (uint32_t)p.tap_shift[tap] * 128u;
          if (!kStream && local == 0 && ph.load_weights) {  // resident weights: tap by tap during the first tile
            ptx::mbar_wait(b_full + 8 * tap, 0);
            ptx::tc_fence_after();
          }
          for (int kb = 0; kb < p.nkb; ++kb) {
            if (kStream) {
              ptx::mbar_wait(b_full + 8 * bs, bphase);
              ptx::tc_fence_after();
              wt = wbuf + bs * wtile_bytes;
            }
            const uint64_t adesc = ptx::make_kmajor_desc(arow + kb * pblock_bytes, 128);
            const uint64_t bdesc = ptx::make_kmajor_desc(wt, 128);
            const int ksteps = min(64, p.cin_pad - kb * 64) >> 4;
            if (kPair) {
              ptx::umma_bf16_pair_elect(tmem_d, adesc, bdesc, idesc, accumulate);
              if (ksteps > 1) ptx::umma_bf16_pair_elect(tmem_d, adesc + 2, bdesc + 2, idesc, 1);
              if (ksteps > 2) ptx::umma_bf16_pair_elect(tmem_d, adesc + 4, bdesc + 4, idesc, 1);
              if (ksteps > 3) ptx::umma_bf16_pair_elect(tmem_d, adesc + 6, bdesc + 6, idesc, 1);
            } else {
              ptx::umma_bf16_elect(tmem_d, adesc, bdesc, idesc, accumulate);
              if (ksteps > 1) ptx::umma_bf16_elect(tmem_d, adesc + 2, bdesc + 2, idesc, 1);
              if (ksteps > 2) ptx::umma_bf16_elect(tmem_d, adesc + 4, bdesc + 4, idesc, 1);
              if (ksteps > 3) ptx::umma_bf16_elect(tmem_d, adesc + 6, bdesc + 6, idesc, 1);
            }
            accumulate = 1;
            if (kStream) {
              if (kPair)
                ptx::umma_commit_pair_elect(b_empty + 8 * bs, 3);
              else
                ptx::umma_commit_elect(b_empty + 8 * bs);
              if (++bs == (uint32_t)p.b_stages) {
                bs = 0;
                bphase ^= 1;
              }
            } else {
              wt += wtile_bytes;
            }
          }
          if (tail_row) {
            // the trailing 16 / 32 channels: one / two K steps from the narrow-swizzle blocks
            const uint64_t adesc = ptx::make_kmajor_desc(
                pbuf + pb * patch_bytes + ptail_off + (uint32_t)p.tap_shift[tap] * tail_row, tail_row);
            const uint64_t bdesc = ptx::make_kmajor_desc(wt, tail_row);
            ptx::umma_bf16_elect(tmem_d, adesc, bdesc, idesc, accumulate);
            if (tail_row > 32) ptx::umma_bf16_elect(tmem_d, adesc + 2, bdesc + 2, idesc, 1);
            accumulate = 1;
            wt += wtail_bytes;
          }
        }
        if (kPair) {
          ptx::umma_commit_pair_elect(p_empty + 8 * pb, 3);
          ptx::umma_commit_pair_elect(tmem_full_bar + 8 * acc, 3);
        } else {
          ptx::umma_commit_elect(p_empty + 8 * pb);
          ptx::umma_commit_elect(tmem_full_bar + 8 * acc);
        }
        CNT_END(3);
        if (++pb == (uint32_t)p.p_bufs) {
          pb = 0;
          pph ^= 1;
        }
        if (kDyn) {
          if (--dyn_left > 0) {
            ++item;
          } else {
            item = take_claim();
            dyn_left = ph.chunk;
          }
        } else {
          item += item_step;
        }
      }
      CNT_FLUSH(1);
      CNT_FLUSH(2);
      CNT_FLUSH(3);
#ifdef RNB_PATCH_COUNTERS
      if (blockIdx.x == 0 && lane == 0) {
        g_patch_cnt[6] += (unsigned long long)(clock64() - loop_c0);
        g_patch_cnt[7] += gtimer() - loop_g0;
      }
#endif
    }
  } else {
    // ------------------------------------------------------------------ epilogue (256 threads)
    ptx::grid_dep_wait();  // residual reads and output stores must not pass the previous kernel
    const int ew = warp - 2;
    const int q = warp & 3;
    const int half = ew >> 2;
    const int et = threadIdx.x - 64;
    const int row = q * 32 + lane;
    const int rw = row & ((1 << p.lpw) - 1);
    const int rh = (row >> p.lpw) & ((1 << p.lph) - 1);
    const int rt = row >> (p.lpw + p.lph);
    const bool in_tile = (rw < p.tw) && (rh < p.th) && (rt < p.tt);
    int local = 0;
    int table_n0 = kStream ? -1 : 0;
#ifdef RNB_PATCH_COUNTERS
    unsigned long long epi_cnt[6] = {};
#endif
    // this thread's slice of a work item: where its output row lives and whether it exists
    struct Slice {
      size_t off;
      int n0, ncols, b;
      bool valid;
    };
    auto slice_of = [&](int item) {
      const PatchItem tc = decode_patch_item(p, item, kNumCta, rank);
      const int w = tc.w0 + rw, h = tc.h0 + rh, t = tc.t0 + rt;
      Slice s;
      s.valid = in_tile && (w < p.Wo) && (h < p.Ho) && (t < p.To) && (tc.b < p.B);
      s.off = ((((size_t)tc.b * p.To + t) * p.Ho + h) * p.Wo + w) * p.cout_pad + tc.n0;
      s.n0 = tc.n0;
      s.ncols = tc.ncols;
      s.b = tc.b;
      return s;
    };
    // One work item.  The residual rows of the item after the next are prefetched into L2 before
    // the wait for this item's accumulator, two tile periods ahead of their use (one period is
    // about one loaded DRAM round trip, which is not enough).
    auto one_item = [&](int item, const Slice& cur, const Slice& nxt) {
      const uint32_t acc = local & 1;
      const uint32_t acc_phase = (local >> 1) & 1;
      Slice nxt2 = nxt;
      if (!kDyn && item + 2 * item_step < total_items) {
        nxt2 = slice_of(item + 2 * item_step);
        if (kRes && !res_smem) prefetch_residual_l2(p.residual + nxt2.off, nxt2.ncols, nxt2.valid);
      }
      if (kStream && cur.n0 != table_n0) {
        // shift table of this channel tile; the barriers order it against the previous readers
        ptx::named_bar_sync(1, kEpiThreads);
        if (et < cur.ncols) {
          const float v = __ldg(p.shift + cur.n0 + et);
          asm volatile("st.shared.f32 [%0], %1;\n" ::"r"(shift_tab + et * 4), "f"(v) : "memory");
        }
        ptx::named_bar_sync(1, kEpiThreads);
        table_n0 = cur.n0;
      }
      CNT_BEGIN();
      const uint32_t taddr_row = tmem_base + ((uint32_t)(q * 32) << 16) + acc * kMaxBlockN;
      epilogue_tile<kRes, kRelu>(tmem_full_bar + 8 * acc, acc_phase, taddr_row, shift_tab,
                                 cur.ncols >> 4, half, cur.valid, p.y + cur.off,
                                 kRes ? p.residual + cur.off : nullptr
#ifdef RNB_PATCH_COUNTERS
                                 ,
                                 epi_cnt
#endif
                                 ,
                                 res_smem ? rbuf + acc * rtile_bytes : 0u, row, res_full + 8 * acc,
                                 acc_phase);
      if (res_smem) {
        // this warp has read its residual rows: the tile buffer may be refilled
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(res_empty + 8 * acc);
      }
      // every TMEM read of this accumulator buffer has completed: hand it back, one arrival per
      // warp (hundreds of arrivals per tile on one mbarrier serialise)
      ptx::tc_fence_before();
      __syncwarp();
      if (kPair) {
        ptx::named_bar_sync(2, kEpiThreads);
        if (et == 0) ptx::mbar_arrive_cluster(tmem_empty_ld + 8 * acc);
      } else {
        if (lane == 0) ptx::mbar_arrive(tmem_empty_bar + 8 * acc);
      }
      if (ph.sig != nullptr) {
        // co-scheduled producer phase: every thread's stores of this item are ordered before the
        // bump of its clip's counter (fence, then the block-level barrier, then one atomic)
        __threadfence();
        ptx::named_bar_sync(3, kEpiThreads);
        if (et == 0 && cur.b < p.B) atomicAdd(ph.sig + cur.b, 1ull);
      }
      CNT_END(5);
      ++local;
      return nxt2;
    };
    Slice cur = {}, nxt = {};
    if (kDyn) {
      // claimed items are only known a claim at a time: no look-ahead (the dynamic phases run
      // layers whose residual comes through shared memory or that have none)
      int item = take_claim();
      dyn_left = ph.chunk;
      while (item < total_items) {
        cur = slice_of(item);
        one_item(item, cur, cur);
        if (--dyn_left > 0) {
          ++item;
        } else {
          item = take_claim();
          dyn_left = ph.chunk;
        }
      }
    } else {
    if (first_item < total_items) {
      cur = slice_of(first_item);
      if (kRes && !res_smem) prefetch_residual_l2(p.residual + cur.off, cur.ncols, cur.valid);
    }
    if (first_item + item_step < total_items) {
      nxt = slice_of(first_item + item_step);
      if (kRes && !res_smem) prefetch_residual_l2(p.residual + nxt.off, nxt.ncols, nxt.valid);
    }
    for (int item = first_item; item < total_items; item += item_step) {
      const Slice nxt2 = one_item(item, cur, nxt);
      cur = nxt;
      nxt = nxt2;
    }
    }
    if (warp == 2) {
      CNT_FLUSH(4);
      CNT_FLUSH(5);
    }
#ifdef RNB_PATCH_COUNTERS
    if (blockIdx.x == 0 && warp == 2 && lane == 0)
      for (int k = 0; k < 6; ++k) g_patch_epi[k] += epi_cnt[k];
#endif
  }

  // end of the phase: every load has landed and been consumed, every MMA has completed (the
  // epilogue waited for the last accumulator), so the next phase may reuse the shared memory
  ptx::tc_fence_before();
  __syncthreads();
  // the leader's last commits may still be on their way to the peer's barriers
  if (kPair) ptx::cluster_sync_all();
  ptx::tc_fence_after();
}

// shared prologue / epilogue of the kernels: TMEM allocation, PDL trigger
template <bool kPair>
__device__ __forceinline__ uint32_t patch_kernel_begin(uint32_t smem_base, int warp) {
  const uint32_t tmem_ptr_smem = smem_base + kPatchBarSlots * kPatchBarSlotBytes + kMaxBlockN * 4u;
  if (warp == 1) {
    if (kPair) {
      ptx::tmem_alloc_pair(tmem_ptr_smem, kTmemCols);
      ptx::tmem_relinquish_pair();
    } else {
      ptx::tmem_alloc(tmem_ptr_smem, kTmemCols);
      ptx::tmem_relinquish();
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  ptx::grid_dep_launch_dependents();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];\n" : "=r"(tmem_base) : "r"(tmem_ptr_smem));
  return __shfl_sync(0xffffffffu, tmem_base, 0);
}

template <bool kRes, bool kRelu, bool kStream, bool kPair>
__global__ void __launch_bounds__(kConvThreads, 1)
conv_patch_kernel(const __grid_constant__ ConvPatchParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
#ifdef RNB_PATCH_COUNTERS
  if (threadIdx.x == 0 && blockIdx.x < 256) {
    g_patch_cta[p.dbg_launch & 3][0][blockIdx.x] = gtimer();
    unsigned int smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    g_patch_smid[p.dbg_launch & 3][blockIdx.x] = smid;
  }
#endif
  // shuffle-broadcast so the compiler knows these are warp-uniform (role branches, TMEM address)
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  // rank inside the 2-CTA cluster; for a (2,1,1) cluster over a 1-D grid that is blockIdx.x & 1,
  // which -- unlike a read of %cluster_ctarank -- the compiler can prove uniform
  const uint32_t rank = kPair ? (blockIdx.x & 1u) : 0u;
  constexpr int kNumCta = kPair ? 2 : 1;
  const uint32_t tmem_base = patch_kernel_begin<kPair>(smem_base, warp);

  const int total_tiles = p.tiles_w * p.tiles_h * p.tiles_t * p.tiles_b;
  PatchPhase ph = {};
  ph.first_item = blockIdx.x / kNumCta;
  ph.item_step = gridDim.x / kNumCta;
  ph.item_lo = 0;
  ph.item_hi = ((total_tiles + kNumCta - 1) / kNumCta) * p.n_tiles;
  ph.load_weights = 1;
  conv_patch_phase<kRes, kRelu, kStream, kPair>(p, ph, smem_base, tmem_base, warp, lane, rank);

  if (warp == 1) {
    if (kPair)
      ptx::tmem_dealloc_pair(tmem_base, kTmemCols);
    else
      ptx::tmem_dealloc(tmem_base, kTmemCols);
#ifdef RNB_PATCH_COUNTERS
    // (the CTA's exit stamp; a stamp taken by thread 0 right after the __syncthreads above was
    // seen to be hoisted over it and showed the PRODUCER's finish time, several us early)
    if (lane == 0 && blockIdx.x < 256) g_patch_cta[p.dbg_launch & 3][2][blockIdx.x] = gtimer();
#endif
  }
}

// ---------------------------------------------------------------------------------------------
// Co-scheduled SpatioTemporalConv: ONE launch runs the 1 x k x k convolution (params `sp`) on the
// first n_sp CTAs and the k x 1 x 1 convolution that consumes its output (params `tp`) on the
// rest, so that the intermediate tensor is read back while it is still in L2 instead of after a
// kernel boundary, when most of it has been evicted to HBM (conv2: 115 MB at batch 16).
//
//   CTA class S (blockIdx < n_sp):  phase 0  all spatial items, round-robin over the S CTAs; each
//                                            finished item bumps its clip's counter in `flags`
//                                   phase 2  joins the temporal work (weights swapped first)
//   CTA class T (the others):       phase 1  temporal items from the start
//
// Temporal items are CLAIMED (PatchPhase::claim), `chunk` at a time, from one grid-wide counter, in
// clip order: the T CTAs work through them as the clips complete, and the S CTAs take their share
// of whatever is left when the spatial work is done -- no static split to get wrong.  A temporal
// item of clip b loads its patch only once flags[b] says every spatial tile of the clip has been
// stored.  Spatial CTAs never wait for temporal ones and all CTAs of the grid are co-resident
// (grid <= SM count, one CTA per SM), so the waits cannot deadlock.  flags[flag_clips] is the
// claim counter, flags[flag_clips + 1] the exit ticket: the last CTA to finish zeroes everything
// for the next launch.
template <bool kResT, bool kReluT>
__global__ void __launch_bounds__(kConvThreads, 1)
conv_patch_dual_kernel(const __grid_constant__ ConvPatchParams sp,
                       const __grid_constant__ ConvPatchParams tp, int n_sp, int chunk,
                       unsigned long long* flags, int flag_clips) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const uint32_t tmem_base = patch_kernel_begin<false>(smem_base, warp);

  const int cta = blockIdx.x, ncta = gridDim.x;
  if (cta < n_sp) {
    PatchPhase ph = {};
    ph.first_item = cta;
    ph.item_step = n_sp;
    ph.item_lo = 0;
    ph.item_hi = sp.tiles_w * sp.tiles_h * sp.tiles_t * sp.tiles_b * sp.n_tiles;
    ph.bar_slot = 0;
    ph.load_weights = 1;
    ph.sig = flags;
    conv_patch_phase<false, true, false, false, false>(sp, ph, smem_base, tmem_base, warp, lane, 0);
  }
  {
    PatchPhase pt = {};
    pt.item_lo = 0;
    pt.item_hi = tp.tiles_w * tp.tiles_h * tp.tiles_t * tp.tiles_b * tp.n_tiles;
    pt.bar_slot = 1;
    pt.load_weights = 1;
    pt.dep = flags;
    pt.dep_target = (unsigned long long)(sp.tiles_w * sp.tiles_h * sp.tiles_t);
    pt.claim = flags + flag_clips;
    pt.chunk = chunk;
    conv_patch_phase<kResT, kReluT, false, false, true>(tp, pt, smem_base, tmem_base, warp, lane, 0);
  }

  if (warp == 1) ptx::tmem_dealloc(tmem_base, kTmemCols);
  if (threadIdx.x == 0) {
    // every CTA is past its last counter access: the last one to get here re-arms the counters
    __threadfence();
    const unsigned long long t = atomicAdd(flags + flag_clips + 1, 1ull);
    if (t == (unsigned long long)ncta - 1) {
      for (int b = 0; b <= flag_clips + 1; ++b) flags[b] = 0ull;
      __threadfence();
    }
  }
}

#ifdef RNB_PATCH_COUNTERS
void patch_cta_times(unsigned long long out[4][3][256]) {
  cudaMemcpyFromSymbol(out, g_patch_cta, sizeof(g_patch_cta));
}
void patch_cta_smids(unsigned int out[4][256]) {
  cudaMemcpyFromSymbol(out, g_patch_smid, sizeof(g_patch_smid));
}
void patch_counters_read_reset(unsigned long long out[8]) {
  cudaMemcpyFromSymbol(out, g_patch_cnt, sizeof(g_patch_cnt));
  unsigned long long z[8] = {};
  cudaMemcpyToSymbol(g_patch_cnt, z, sizeof(z));
}
void patch_epi_read_reset(unsigned long long out[8]) {
  cudaMemcpyFromSymbol(out, g_patch_epi, sizeof(g_patch_epi));
  unsigned long long z[8] = {};
  cudaMemcpyToSymbol(g_patch_epi, z, sizeof(z));
}
#endif

namespace {
template <bool kRes, bool kRelu, bool kStream, bool kPair>
cudaError_t launch_patch_variant(const ConvPatchParams& p, int num_sms, int smem,
                                 cudaStream_t stream) {
  static std::atomic<bool> configured[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    e = cudaFuncSetAttribute(conv_patch_kernel<kRes, kRelu, kStream, kPair>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit);
    if (e != cudaSuccess) return e;
    configured[dev] = true;
  }
  const int ncta = kPair ? 2 : 1;
  const int tiles = p.tiles_w * p.tiles_h * p.tiles_t * p.tiles_b;
  const int items = ((tiles + ncta - 1) / ncta) * p.n_tiles;
  int max_groups = num_sms / ncta;
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(kConvThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[3];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (kPair) {
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = 2;
    attr[1].val.clusterDim.y = 1;
    attr[1].val.clusterDim.z = 1;
    cfg.numAttrs = 2;
    // A persistent grid must be fully resident.  Not every SM has a partner for a 2-CTA cluster
    // (SMs fused off leave single-SM TPCs), so ask how many pairs fit at once.
    static std::atomic<int> cached[64] = {};
    if (dev >= 0 && dev < 64 && cached[dev] == 0) {
      cfg.gridDim = dim3(2 * (num_sms / 2));
      int n = 0;
      e = cudaOccupancyMaxActiveClusters(&n, conv_patch_kernel<kRes, kRelu, kStream, kPair>, &cfg);
      if (e != cudaSuccess) return e;
      cached[dev] = n > 0 ? n : 1;
    }
    if (dev >= 0 && dev < 64 && cached[dev] < max_groups) max_groups = cached[dev];
  }
  const int groups = items < max_groups ? items : max_groups;
  cfg.gridDim = dim3(groups * ncta);
  cfg.numAttrs = add_l2_output_window(attr, cfg.numAttrs, p.y,
                                      (size_t)p.B * p.To * p.Ho * p.Wo * p.cout_pad * 2);
  return cudaLaunchKernelEx(&cfg, conv_patch_kernel<kRes, kRelu, kStream, kPair>, p);
}

template <bool kStream, bool kPair>
cudaError_t launch_patch_mode(const ConvPatchParams& p, int num_sms, int smem, cudaStream_t stream) {
  const bool res = p.residual != nullptr, relu = p.relu != 0;
  if (res)
    return relu ? launch_patch_variant<true, true, kStream, kPair>(p, num_sms, smem, stream)
                : launch_patch_variant<true, false, kStream, kPair>(p, num_sms, smem, stream);
  return relu ? launch_patch_variant<false, true, kStream, kPair>(p, num_sms, smem, stream)
              : launch_patch_variant<false, false, kStream, kPair>(p, num_sms, smem, stream);
}
}  // namespace

namespace {
int patch_params_smem(const ConvPatchParams& p) {
  const int ncta = p.pair ? 2 : 1;
  return conv_patch_smem_bytes(p.nkb, p.patch_rows, p.num_taps, p.block_n / ncta, p.b_stages,
                               p.p_bufs, p.tail_ch, (p.res_smem && p.residual) ? p.n / 64 : 0);
}

template <bool kResT, bool kReluT>
cudaError_t launch_dual_variant(const ConvPatchParams& sp, const ConvPatchParams& tp, int n_sp,
                                int chunk, unsigned long long* flags, int flag_clips, int grid,
                                int smem, cudaStream_t stream) {
  static std::atomic<bool> configured[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    e = cudaFuncSetAttribute(conv_patch_dual_kernel<kResT, kReluT>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit);
    if (e != cudaSuccess) return e;
    configured[dev] = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kConvThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[3];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  // the intermediate tensor is what should stay in L2 for this launch
  cfg.numAttrs = add_l2_output_window(attr, cfg.numAttrs, sp.y,
                                      (size_t)sp.B * sp.To * sp.Ho * sp.Wo * sp.cout_pad * 2);
  return cudaLaunchKernelEx(&cfg, conv_patch_dual_kernel<kResT, kReluT>, sp, tp, n_sp, chunk,
                            flags, flag_clips);
}
}  // namespace

bool conv_patch_dual_applicable(const ConvPatchParams& sp, const ConvPatchParams& tp) {
  // resident weights, one CTA per tile, one channel tile; the temporal patches are cut from the
  // spatial output of the same clips
  return sp.b_stages == 0 && tp.b_stages == 0 && !sp.pair && !tp.pair && sp.n_tiles == 1 &&
         tp.n_tiles == 1 && sp.B == tp.B && sp.tiles_b == sp.B && tp.tiles_b == tp.B &&
         sp.residual == nullptr && sp.relu != 0 && (tp.residual == nullptr || tp.res_smem);
}

cudaError_t launch_conv_patch_dual(const ConvPatchParams& sp, const ConvPatchParams& tp, int n_sp,
                                   int chunk, unsigned long long* flags, int flag_clips,
                                   int num_sms, cudaStream_t stream) {
  if (!conv_patch_dual_applicable(sp, tp) || !flags || flag_clips < sp.B) return cudaErrorInvalidValue;
  const int grid = num_sms;
  if (n_sp < 1 || n_sp >= grid || chunk < 1) return cudaErrorInvalidValue;
  const int a = patch_params_smem(sp), b = patch_params_smem(tp);
  const int smem = a > b ? a : b;
  if (smem > kSmemLimit) return cudaErrorInvalidValue;
  const bool res = tp.residual != nullptr, relu = tp.relu != 0;
  if (res)
    return relu ? launch_dual_variant<true, true>(sp, tp, n_sp, chunk, flags, flag_clips, grid, smem,
                                                  stream)
                : launch_dual_variant<true, false>(sp, tp, n_sp, chunk, flags, flag_clips, grid,
                                                   smem, stream);
  return relu ? launch_dual_variant<false, true>(sp, tp, n_sp, chunk, flags, flag_clips, grid, smem,
                                                 stream)
              : launch_dual_variant<false, false>(sp, tp, n_sp, chunk, flags, flag_clips, grid, smem,
                                                  stream);
}

cudaError_t launch_conv_patch(const ConvPatchParams& p, int num_sms, cudaStream_t stream) {
  if (p.b_stages < 0 || p.b_stages > kPatchMaxBStages) return cudaErrorInvalidValue;
  const int ncta = p.pair ? 2 : 1;
  if (p.p_bufs < 2 || p.p_bufs > kPatchMaxBufs) return cudaErrorInvalidValue;
  const int smem = patch_params_smem(p);
  (void)ncta;
  if (p.pair)
    return p.b_stages > 0 ? launch_patch_mode<true, true>(p, num_sms, smem, stream)
                          : launch_patch_mode<false, true>(p, num_sms, smem, stream);
  return p.b_stages > 0 ? launch_patch_mode<true, false>(p, num_sms, smem, stream)
                        : launch_patch_mode<false, false>(p, num_sms, smem, stream);
}

}  // namespace rnb
