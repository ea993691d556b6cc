#include "plan.cuh"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

namespace rnb {

uint16_t f32_to_bf16_bits(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);  // NaN
  const uint32_t lsb = (u >> 16) & 1u;
  u += 0x7fffu + lsb;  // round to nearest even
  return (uint16_t)(u >> 16);
}

float bf16_bits_to_f32(uint16_t b) {
  const uint32_t u = (uint32_t)b << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

void pack_conv_weights_tc(const ConvGeom& g, const float* w, const float* scale,
                          std::vector<uint16_t>* packed, const float* w_extra,
                          const float* scale_extra, const float* w_ncat, const float* scale_ncat) {
  const int kp = g.k_total();
  const int cp = g.cin_pad();
  packed->assign((size_t)g.cout_pad() * kp, 0);
  for (int n = 0; n < g.cout; ++n) {
    uint16_t* row = packed->data() + (size_t)n * kp;
    const float sc = scale ? scale[n] : 1.f;
    for (int c = 0; c < g.cin; ++c)
      for (int dt = 0; dt < g.kt; ++dt)
        for (int dh = 0; dh < g.kh; ++dh)
          for (int dw = 0; dw < g.kw; ++dw) {
            const float v = w[((((size_t)n * g.cin + c) * g.kt + dt) * g.kh + dh) * g.kw + dw];
            size_t k;
            if (g.stem_unfolded) {
              // filter row dh reads input row 2*ho + dh - 3 = 2*(ho + j - 2) + par: row pair -> tap,
              // (row of the pair, column, channel) -> K
              const int off = dh - 3;
              const int par = ((off % 2) + 2) % 2;
              const int j = (off - par) / 2 + 2;
              k = (size_t)j * cp + par * 32 + dw * 4 + c;
            } else {
              k = (size_t)((dt * g.kh + dh) * g.kw + dw) * cp + c;
            }
            row[k] = f32_to_bf16_bits(v * sc);
          }
    if (g.extra_cin && w_extra) {
      const float se = scale_extra ? scale_extra[n] : 1.f;
      for (int c = 0; c < g.extra_cin; ++c)
        row[(size_t)g.num_taps() * cp + c] =
            f32_to_bf16_bits(w_extra[(size_t)n * g.extra_cin + c] * se);
    }
  }
  if (g.ncat && w_ncat) {
    // rows pad16(cout) ..: non-zero only in the K slice of the tap at offset zero
    const int centre = (g.pt * g.kh + g.ph) * g.kw + g.pw;
    for (int n = 0; n < g.ncat; ++n) {
      uint16_t* row = packed->data() + (size_t)(pad16(g.cout) + n) * kp;
      const float sc = scale_ncat ? scale_ncat[n] : 1.f;
      for (int c = 0; c < g.cin; ++c)
        row[(size_t)centre * cp + c] = f32_to_bf16_bits(w_ncat[(size_t)n * g.cin + c] * sc);
    }
  }
}

void pack_conv_weights_ref(const ConvGeom& g, const float* w, std::vector<float>* packed) {
  packed->assign((size_t)g.kt * g.kh * g.kw * g.cin * g.cout, 0.f);
  for (int n = 0; n < g.cout; ++n)
    for (int c = 0; c < g.cin; ++c)
      for (int dt = 0; dt < g.kt; ++dt)
        for (int dh = 0; dh < g.kh; ++dh)
          for (int dw = 0; dw < g.kw; ++dw)
            (*packed)[((((size_t)dt * g.kh + dh) * g.kw + dw) * g.cin + c) * g.cout + n] =
                w[((((size_t)n * g.cin + c) * g.kt + dt) * g.kh + dh) * g.kw + dw];
}

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn(std::string* err) {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  static std::string once_err;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || p == nullptr) {
      once_err = std::string("cuTensorMapEncodeTiled not available: ") + cudaGetErrorString(e);
      (void)cudaGetLastError();
    } else {
      fn = reinterpret_cast<EncodeTiledFn>(p);
    }
  });
  if (!fn && err) *err = once_err;
  return fn;
}

// bf16 tensor map of rank `rank`; dims/box innermost first; strides in BYTES for dims 1..rank-1
bool encode_bf16(CUtensorMap* m, const void* base, int rank, const uint64_t* dims,
                 const uint64_t* strides, const uint32_t* box, int swizzle_bytes, std::string* err) {
  EncodeTiledFn fn = get_encode_fn(err);
  if (!fn) return false;
  cuuint64_t gdim[5], gstr[4];
  cuuint32_t bdim[5], estr[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bdim[i] = box[i];
    estr[i] = 1;
  }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides[i];
  const CUtensorMapSwizzle swz = swizzle_bytes == 128  ? CU_TENSOR_MAP_SWIZZLE_128B
                                 : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                                       : CU_TENSOR_MAP_SWIZZLE_32B;
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base),
                  gdim, gstr, bdim, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    if (err) {
      char buf[512];
      snprintf(buf, sizeof buf,
               "cuTensorMapEncodeTiled failed (%d): rank %d dims [%llu %llu %llu %llu %llu] box "
               "[%u %u %u %u %u] stride0 %llu swizzle %d base %p",
               (int)r, rank, (unsigned long long)gdim[0], (unsigned long long)(rank > 1 ? gdim[1] : 0),
               (unsigned long long)(rank > 2 ? gdim[2] : 0), (unsigned long long)(rank > 3 ? gdim[3] : 0),
               (unsigned long long)(rank > 4 ? gdim[4] : 0), bdim[0], rank > 1 ? bdim[1] : 0,
               rank > 2 ? bdim[2] : 0, rank > 3 ? bdim[3] : 0, rank > 4 ? bdim[4] : 0,
               (unsigned long long)(rank > 1 ? gstr[0] : 0), swizzle_bytes, base);
      *err = buf;
    }
    return false;
  }
  return true;
}

int ceil_div(int a, int b) { return (a + b - 1) / b; }

// position box (bw, bh, bt, bb), powers of two with product 128, minimising the number of tiles
void choose_box(int Wo, int Ho, int To, int B, int* lbw, int* lbh, int* lbt) {
  long long best = -1;
  int bw_ = 0, bh_ = 0, bt_ = 0;
  for (int a = 0; a <= 7; ++a)
    for (int b = 0; a + b <= 7; ++b)
      for (int c = 0; a + b + c <= 7; ++c) {
        const int d = 7 - a - b - c;
        const long long tiles = (long long)ceil_div(Wo, 1 << a) * ceil_div(Ho, 1 << b) *
                                ceil_div(To, 1 << c) * ceil_div(B, 1 << d);
        // fewer tiles first; then longer contiguous runs (w, then h, then t)
        const long long key = tiles * 4096 - (a * 64 + b * 8 + c);
        if (best < 0 || key < best) {
          best = key;
          bw_ = a;
          bh_ = b;
          bt_ = c;
        }
      }
  *lbw = bw_;
  *lbh = bh_;
  *lbt = bt_;
}

struct View {
  int pw, ph, pt;
};

}  // namespace

bool build_conv_tc_params(const ConvGeom& g, int B, const __nv_bfloat16* x,
                          const __nv_bfloat16* w_packed, __nv_bfloat16* y, const float* shift,
                          const __nv_bfloat16* residual, bool relu, int num_sms, ConvTcParams* out,
                          std::string* err, int force_cm, int force_cn, int split_k, int split_n,
                          int pair, const __nv_bfloat16* x_extra) {
  ConvTcParams& p = *out;
  memset(&p, 0, sizeof p);
  const int cp = g.cin_pad();
  const int np = g.cout_pad();
  for (int s : {g.st, g.sh, g.sw})
    if (s != 1 && s != 2) {
      if (err) *err = "conv_tc: only strides 1 and 2 are supported";
      return false;
    }
  if ((g.sw == 2 && (g.Wi & 1)) || (g.sh == 2 && (g.Hi & 1)) || (g.st == 2 && (g.Ti & 1))) {
    if (err) *err = "conv_tc: a strided dimension must have even extent";
    return false;
  }
  p.shift = shift;
  p.residual = residual;
  p.relu = relu ? 1 : 0;
  p.cin_pad = cp;
  if (g.stem_unfolded) {
    if (err) *err = "conv_tc: the unfolded stem runs through the patch kernel";
    return false;
  }
  p.bk = 64;
  p.kb_per_tap = ceil_div(cp, p.bk);
  p.Wo = g.Wo;
  p.Ho = g.Ho;
  p.To = g.To;
  p.B = B;
  p.cout_pad = np;

  choose_box(g.Wo, g.Ho, g.To, B, &p.lbw, &p.lbh, &p.lbt);
  const int bw = 1 << p.lbw, bh = 1 << p.lbh, bt = 1 << p.lbt, bb = 128 >> (p.lbw + p.lbh + p.lbt);
  p.tiles_w = ceil_div(g.Wo, bw);
  p.tiles_h = ceil_div(g.Ho, bh);
  p.tiles_t = ceil_div(g.To, bt);
  p.tiles_b = ceil_div(B, bb);
  const int m_tiles = p.tiles_w * p.tiles_h * p.tiles_t * p.tiles_b;

  // channel tiling: one tile if it fits the accumulator, otherwise multiples of 64 channels so
  // that the 64-channel output sub-tiles of neighbouring tiles never overlap
  if (np <= kMaxBlockN) {
    p.block_n = np;
    p.n_tiles = 1;
  } else {
    const int nt = ceil_div(np, kMaxBlockN);
    p.block_n = round_up(ceil_div(np, nt), 64);
    p.n_tiles = ceil_div(np, p.block_n);
  }
  if (g.extra_cin && (split_k > 1 || !x_extra)) {
    if (err) *err = "conv_tc: a fused shortcut tap needs its tensor and excludes split-K";
    return false;
  }
  if (pair && (split_k > 1 || m_tiles < 2)) {
    if (err) *err = "conv_tc: CTA pairs need two position tiles and exclude split-K";
    return false;
  }
  if (split_k > 1 || pair) {
    // split-K (conv_splitk.cu): keep the channel tiles as wide as the accumulator allows -- the
    // parallelism comes from K -- in near-equal multiples of 16 channels.  CTA pairs
    // (conv_pair.cu): wide tiles too, each CTA holds half of one.
    const int cap = split_n >= 16 && split_n < kMaxBlockN ? split_n : kMaxBlockN;
    const int nt = ceil_div(np, cap);
    p.block_n = round_up(ceil_div(np, nt), 16);
    p.n_tiles = ceil_div(np, p.block_n);
    const int total_kb = g.kt * g.kh * g.kw * p.kb_per_tap;  // upper bound; re-checked below
    if (split_k > 1 && (split_k > kMaxSplitK || split_k > total_kb)) {
      if (err) *err = "conv_tc: split-K factor too large";
      return false;
    }
  } else if (split_n >= 16) {
    // the caller picks the channel tile (e.g. so that position tiles x channel tiles fill one wave)
    p.block_n = round_up(split_n, 16) < kMaxBlockN ? round_up(split_n, 16) : kMaxBlockN;
    p.n_tiles = ceil_div(np, p.block_n);
  } else {
    // small layers: more, narrower tiles until (nearly) every SM has one -- 147 wide tiles on
    // 148 SMs are one full wave and must not be cut into 245 narrow ones (1.65 waves)
    while (m_tiles * p.n_tiles * 10 < num_sms * 9 && p.block_n > 64) {
      p.block_n = round_up(p.block_n, 64) - 64;
      p.n_tiles = ceil_div(np, p.block_n);
    }
  }
  const int stage_bytes =
      pair ? conv_pair_stage_bytes(p.bk, p.block_n) : conv_tc_stage_bytes(p.bk, p.block_n);
  int stages = (kSmemLimit - conv_tc_smem_bytes(p.bk, p.block_n, 0)) / stage_bytes;
  if (stages > 10) stages = 10;
  if (stages < 2) {
    if (err) *err = "conv_tc: pipeline does not fit in shared memory";
    return false;
  }
  p.num_stages = stages;

  // ---- thread-block cluster: share operand tiles between CTAs by TMA multicast where that cuts
  // the L2 -> shared-memory traffic per tile (the bound of the conv3-5 layers)
  p.m_tiles = m_tiles;
  // the activation box is split along its slowest non-unit dimension so slices stay contiguous
  const int ext[4] = {bw, bh, bt, bb};
  int sdim = 0;
  for (int d = 3; d >= 0; --d)
    if (ext[d] > 1) {
      sdim = d;
      break;
    }
  int best_cm = 1, best_cn = 1;
  {
    if (force_cm <= 0)
      if (const char* e = getenv("RNB_B200_CLUSTER")) sscanf(e, "%dx%d", &force_cm, &force_cn);
    const double a_b = (double)kTileM * p.bk * 2;
    const double b_b = (double)np / p.n_tiles * p.bk * 2;
    double best_cost = -1;
    static const int cand[6][2] = {{1, 1}, {2, 1}, {1, 2}, {2, 2}, {4, 1}, {1, 4}};
    for (const auto& c : cand) {
      const int cm = c[0], cn = c[1], cs = cm * cn;
      if (force_cm > 0 && (cm != force_cm || cn != force_cn)) continue;
      if (cn > 1 && ext[sdim] % cn != 0) continue;
      if (cm > 1 && (p.block_n % (8 * cm) != 0)) continue;
      // clusters of 4 cannot use every SM (GPC boundaries); measured ~33 of 37 on B200
      const int max_clusters = cs == 4 ? (num_sms / 4) * 9 / 10 : num_sms / cs;
      const int items = ceil_div(m_tiles, cm) * ceil_div(p.n_tiles, cn);
      const int rounds = ceil_div(items, max_clusters > 0 ? max_clusters : 1);
      const double cost = rounds * (a_b / cn + b_b / cm);
      if (best_cost < 0 || cost < best_cost * 0.97) {  // prefer the smaller cluster on near-ties
        best_cost = cost;
        best_cm = cm;
        best_cn = cn;
      }
    }
  }
  if (split_k > 1 || pair) best_cm = best_cn = 1;
  p.ks = split_k > 1 ? split_k : 0;
  p.pair = pair ? 1 : 0;
  p.cm = best_cm;
  p.cn = best_cn;
  p.a_slice_dim = sdim;
  p.a_slice_step = ext[sdim] / p.cn;

  // ---- taps and activation views
  std::vector<View> views;
  auto view_index = [&](int pw, int ph, int pt) {
    for (size_t i = 0; i < views.size(); ++i)
      if (views[i].pw == pw && views[i].ph == ph && views[i].pt == pt) return (int)i;
    views.push_back({pw, ph, pt});
    return (int)views.size() - 1;
  };
  auto split = [](int off, int stride, int* par, int* shift) {
    if (stride == 1) {
      *par = 0;
      *shift = off;
    } else {
      *par = ((off % 2) + 2) % 2;
      *shift = (off - *par) / 2;
    }
  };
  // a tap whose input coordinate lies in the zero padding for EVERY output position contributes
  // nothing (conv5's temporal convolutions run on T = 1: only the centre tap is live)
  auto tap_live = [](int d, int pad, int stride, int in_ext, int out_ext) {
    for (int o = 0; o < out_ext; ++o) {
      const int i = o * stride + d - pad;
      if (i >= 0 && i < in_ext) return true;
    }
    return false;
  };
  int ntap = 0, tap_index = -1;
  for (int dt = 0; dt < g.kt; ++dt)
    for (int dh = 0; dh < g.kh; ++dh)
      for (int dw = 0; dw < g.kw; ++dw) {
        ++tap_index;
        if (!tap_live(dt, g.pt, g.st, g.Ti, g.To) || !tap_live(dh, g.ph, g.sh, g.Hi, g.Ho) ||
            !tap_live(dw, g.pw, g.sw, g.Wi, g.Wo))
          continue;
        if (ntap >= kMaxTaps) {
          if (err) *err = "conv_tc: too many filter taps";
          return false;
        }
        int pw = 0, ph = 0, pt = 0, sw_ = 0, sh_ = 0, st_ = 0;
        split(dt - g.pt, g.st, &pt, &st_);
        split(dh - g.ph, g.sh, &ph, &sh_);
        split(dw - g.pw, g.sw, &pw, &sw_);
        p.tap_map[ntap] = (int8_t)view_index(pw, ph, pt);
        p.tap_dw[ntap] = (int8_t)sw_;
        p.tap_dh[ntap] = (int8_t)sh_;
        p.tap_dt[ntap] = (int8_t)st_;
        p.tap_coff[ntap] = 0;
        p.tap_koff[ntap] = tap_index * cp;
        p.tap_cin[ntap] = (int16_t)cp;
        p.tap_nkb[ntap] = (int8_t)p.kb_per_tap;
        ++ntap;
      }
  int extra_view = -1;
  if (g.extra_cin) {
    // the fused shortcut convolution: one tap at the output position itself, from its own view
    if (ntap >= kMaxTaps || views.size() >= 4) {
      if (err) *err = "conv_tc: no room for the fused shortcut tap";
      return false;
    }
    extra_view = (int)views.size();
    p.tap_map[ntap] = (int8_t)extra_view;
    p.tap_dw[ntap] = p.tap_dh[ntap] = p.tap_dt[ntap] = 0;
    p.tap_coff[ntap] = (int16_t)g.extra_coff;
    p.tap_koff[ntap] = g.num_taps() * cp;
    p.tap_cin[ntap] = (int16_t)g.extra_cin_pad();
    p.tap_nkb[ntap] = (int8_t)ceil_div(g.extra_cin_pad(), p.bk);
    ++ntap;
  }
  p.num_taps = ntap;
  if (p.ks > 1) {
    const int total_kb = ntap * p.kb_per_tap;
    const int per = ceil_div(total_kb, p.ks);
    if ((p.ks - 1) * per >= total_kb) {
      if (err) *err = "conv_tc: split-K factor leaves a CTA without K blocks";
      return false;
    }
  }
  if (views.size() + (extra_view >= 0 ? 1 : 0) > 4) {
    if (err) *err = "conv_tc: more than four strided input views";
    return false;
  }

  // ---- tensor maps
  const int swz = p.bk * 2;
  // input row pitch: the unfolded stem stores Wo groups of 32 per image row
  const uint64_t in_w = (uint64_t)g.Wi;
  const uint64_t in_c = (uint64_t)(g.in_pitch ? g.in_pitch : cp);  // memory pitch of a position
  const int eff_sw = g.sw;
  for (size_t v = 0; v < views.size(); ++v) {
    const View& vw = views[v];
    const __nv_bfloat16* base =
        x + ((size_t)vw.pt * g.Hi * in_w + (size_t)vw.ph * in_w + (size_t)vw.pw) * in_c;
    const uint64_t dims[5] = {in_c, (uint64_t)((in_w - vw.pw + eff_sw - 1) / eff_sw),
                              (uint64_t)((g.Hi - vw.ph + g.sh - 1) / g.sh),
                              (uint64_t)((g.Ti - vw.pt + g.st - 1) / g.st), (uint64_t)B};
    const uint64_t strides[4] = {(uint64_t)eff_sw * in_c * 2, (uint64_t)g.sh * in_w * in_c * 2,
                                 (uint64_t)g.st * g.Hi * in_w * in_c * 2,
                                 (uint64_t)g.Ti * g.Hi * in_w * in_c * 2};
    uint32_t box[5] = {(uint32_t)p.bk, (uint32_t)bw, (uint32_t)bh, (uint32_t)bt, (uint32_t)bb};
    box[1 + p.a_slice_dim] /= (uint32_t)p.cn;  // every CTA of a cluster row fetches one slice
    if (!encode_bf16(&p.a_map[v], base, 5, dims, strides, box, swz, err)) return false;
  }
  for (size_t v = views.size(); v < 4; ++v) p.a_map[v] = p.a_map[0];
  if (extra_view >= 0) {
    // (B, extra_Ti / extra_st, Ho, Wo, extra_pitch): every extra_st-th frame of the second tensor
    const uint64_t pc = (uint64_t)g.extra_pitch;
    const uint64_t dims[5] = {pc, (uint64_t)g.Wo, (uint64_t)g.Ho,
                              (uint64_t)((g.extra_Ti + g.extra_st - 1) / g.extra_st), (uint64_t)B};
    const uint64_t strides[4] = {pc * 2, (uint64_t)g.Wo * pc * 2,
                                 (uint64_t)g.extra_st * g.Ho * g.Wo * pc * 2,
                                 (uint64_t)g.extra_Ti * g.Ho * g.Wo * pc * 2};
    uint32_t box[5] = {(uint32_t)p.bk, (uint32_t)bw, (uint32_t)bh, (uint32_t)bt, (uint32_t)bb};
    box[1 + p.a_slice_dim] /= (uint32_t)p.cn;
    if (!encode_bf16(&p.a_map[extra_view], x_extra, 5, dims, strides, box, swz, err)) return false;
  }
  {
    const uint64_t dims[2] = {(uint64_t)g.k_total(), (uint64_t)np};
    const uint64_t strides[1] = {(uint64_t)g.k_total() * 2};
    // rows one CTA fetches: its share of a multicast tile, or its half of a CTA pair's tile
    const uint32_t box[2] = {(uint32_t)p.bk, (uint32_t)(p.pair ? p.block_n / 2 : p.block_n / p.cm)};
    if (!encode_bf16(&p.b_map, w_packed, 2, dims, strides, box, swz, err)) return false;
  }
  p.y = y;
  return true;
}

namespace {
int ilog2_tp(int v) {
  int l = 0;
  while ((1 << l) < v) ++l;
  return l;
}
}  // namespace

bool conv_tpatch_applicable(const ConvGeom& g) {
  return !g.stem_unfolded && g.kt == 3 && g.kh == 1 && g.kw == 1 && g.st == 1 && g.sh == 1 &&
         g.sw == 1 && g.pt == 1 && g.ph == 0 && g.pw == 0 && g.To >= 2 && g.To == g.Ti &&
         !g.extra_cin && !g.ncat;
}

bool build_conv_tpatch_params(const ConvGeom& g, int B, const __nv_bfloat16* x,
                              const __nv_bfloat16* w_packed, __nv_bfloat16* y, const float* shift,
                              const __nv_bfloat16* residual, bool relu, int n_cap, ConvTcParams* out,
                              std::string* err) {
  if (!conv_tpatch_applicable(g)) {
    if (err) *err = "conv_tpatch: geometry not supported";
    return false;
  }
  ConvTcParams& p = *out;
  memset(&p, 0, sizeof p);
  const int cp = g.cin_pad(), np = g.cout_pad();
  const int hw = g.Ho * g.Wo;
  // tile: tt frames of pt consecutive positions of the flattened H x W plane (a 3x1x1 filter does
  // not care about the spatial structure, and 28 x 28 or 14 x 14 planes tile badly as boxes)
  const int tt = g.To >= 8 ? 8 : (g.To >= 4 ? 4 : 2);
  const int pt = 128 / tt;
  p.shift = shift;
  p.residual = residual;
  p.y = y;
  p.relu = relu ? 1 : 0;
  p.cin_pad = cp;
  p.bk = 64;
  p.kb_per_tap = ceil_div(cp, 64);
  p.num_taps = 3;
  p.Wo = hw;
  p.Ho = 1;
  p.To = g.To;
  p.B = B;
  p.cout_pad = np;
  p.lbw = ilog2_tp(pt);
  p.lbh = 0;
  p.lbt = ilog2_tp(tt);
  p.tiles_w = ceil_div(hw, pt);
  p.tiles_h = 1;
  p.tiles_t = ceil_div(g.To, tt);
  p.tiles_b = B;
  p.m_tiles = p.tiles_w * p.tiles_t * p.tiles_b;
  if (p.m_tiles < 2) {
    if (err) *err = "conv_tpatch: CTA pairs need two position tiles";
    return false;
  }
  const int cap = n_cap >= 16 && n_cap < kMaxBlockN ? n_cap : kMaxBlockN;
  const int nt = ceil_div(np, cap);
  p.block_n = round_up(ceil_div(np, nt), 16);
  p.n_tiles = ceil_div(np, p.block_n);
  p.cm = p.cn = 1;
  p.pair = 1;
  p.tpatch = 1;
  p.tp_rows = (tt + 2) * pt;
  p.tp_shift = pt;
  const int stage = conv_tpatch_stage_bytes(p.tp_rows, p.block_n);
  int stages = (kSmemLimit - 1024 - 256 - kMaxBlockN * 4) / stage;
  if (stages > 10) stages = 10;
  if (stages < 2) {
    if (err) *err = "conv_tpatch: pipeline does not fit in shared memory";
    return false;
  }
  p.num_stages = stages;
  for (int tap = 0; tap < 3; ++tap) {
    p.tap_koff[tap] = tap * cp;
    p.tap_cin[tap] = (int16_t)cp;
    p.tap_nkb[tap] = (int8_t)p.kb_per_tap;
  }
  {
    const uint64_t pc = (uint64_t)(g.in_pitch ? g.in_pitch : cp);
    const uint64_t dims[5] = {pc, (uint64_t)hw, 1, (uint64_t)g.Ti, (uint64_t)B};
    const uint64_t strides[4] = {pc * 2, (uint64_t)hw * pc * 2, (uint64_t)hw * pc * 2,
                                 (uint64_t)g.Ti * hw * pc * 2};
    const uint32_t box[5] = {64u, (uint32_t)pt, 1u, (uint32_t)(tt + 2), 1u};
    if (!encode_bf16(&p.a_map[0], x, 5, dims, strides, box, 128, err)) return false;
    for (int v = 1; v < 4; ++v) p.a_map[v] = p.a_map[0];
  }
  {
    const uint64_t dims[2] = {(uint64_t)g.k_total(), (uint64_t)np};
    const uint64_t strides[1] = {(uint64_t)g.k_total() * 2};
    const uint32_t box[2] = {64u, (uint32_t)(p.block_n / 2)};
    if (!encode_bf16(&p.b_map, w_packed, 2, dims, strides, box, 128, err)) return false;
  }
  return true;
}

namespace {
struct PatchShape {
  bool ok = false;
  bool spatial = false;
  int nkb = 0, patch_rows = 0, taps = 0;
  int tw = 0, th = 0, tt = 0, lpw = 0, lph = 0;
  int pbw = 0, pbh = 0, pbt = 0;  // patch box extent
  int n_tiles = 1, block_n = 0, b_stages = 0, p_bufs = 2;
  int tail_ch = 0;  // trailing 16 / 32 channels in a narrow block of their own (resident mode)
  double eff = 0;
};

int ilog2(int v) {
  int l = 0;
  while ((1 << l) < v) ++l;
  return l;
}

PatchShape patch_shape(const ConvGeom& g, int pair) {
  const int ncta = pair ? 2 : 1;
  PatchShape s;
  // fused shortcuts and channel-sliced inputs: tap / CTA-pair kernel only
  if (g.extra_cin || g.ncat || (g.in_pitch && g.in_pitch != g.cin_pad())) return s;
  if (!g.stem_unfolded && (g.st != 1 || g.sh != 1 || g.sw != 1)) return s;
  s.nkb = ceil_div(g.cin_pad(), 64);
  if (s.nkb > kPatchMaxKb) return s;
  const int np = g.cout_pad();
  if (g.stem_unfolded) {
    // unfolded double rows: 8 output columns x 16 output rows per tile, three halo row pairs
    // (two above, one below); a tap is an offset of one patch row pair = 8 rows
    s.spatial = true;
    s.taps = 4;
    s.tw = 8; s.th = 16; s.tt = 1; s.lpw = 3; s.lph = 4;
    s.pbw = 8; s.pbh = 16 + 3; s.pbt = 1;
    s.eff = (double)g.Wo * g.Ho / ((double)ceil_div(g.Wo, 8) * ceil_div(g.Ho, 16) * 128);
  } else if (g.kt == 1 && g.kh == 3 && g.kw == 3 && g.ph == 1 && g.pw == 1 && g.pt == 0) {
    // One frame per tile.  Accumulator rows use a power-of-two pitch that also holds the two halo
    // columns, so that a filter tap is a pure row offset (dh * pitch + dw) inside the patch.
    s.spatial = true;
    s.taps = 9;
    for (int pitch : {16, 32}) {
      const int th = 128 / pitch;
      const int tiles_w = ceil_div(g.Wo, pitch - 2);
      const int tw = ceil_div(g.Wo, tiles_w);
      const double eff = (double)g.Wo * g.Ho / ((double)tiles_w * ceil_div(g.Ho, th) * 128);
      if (eff > s.eff) {
        s.eff = eff;
        s.tw = tw; s.th = th; s.tt = 1; s.lpw = ilog2(pitch); s.lph = ilog2(th);
        s.pbw = pitch; s.pbh = th + 2; s.pbt = 1;
      }
    }
  } else if (g.kt == 3 && g.kh == 1 && g.kw == 1 && g.pt == 1 && g.ph == 0 && g.pw == 0 &&
             g.To >= 2) {
    // tt frames of P = 128 / tt positions; a tap is an offset of P rows in the (tt + 2)-frame patch
    s.taps = 3;
    const int tt = g.To >= 8 ? 8 : (g.To >= 4 ? 4 : 2);
    const int P = 128 / tt;
    for (int pw = 8; pw <= P && pw <= 64; pw *= 2) {
      const int ph = P / pw;
      const int tiles_w = ceil_div(g.Wo, pw);
      const int tw = ceil_div(g.Wo, tiles_w);
      const int tiles_h = ceil_div(g.Ho, ph);
      const double eff = (double)g.Wo * g.Ho * g.To /
                         ((double)tiles_w * tiles_h * ceil_div(g.To, tt) * 128);
      if (eff > s.eff) {
        s.eff = eff;
        s.tw = tw; s.th = ph; s.tt = tt; s.lpw = ilog2(pw); s.lph = ilog2(ph);
        s.pbw = pw; s.pbh = ph; s.pbt = tt + 2;
      }
    }
  } else {
    return s;
  }
  // ragged tiles are legal (stores are masked) but waste tensor work; require a reasonable fit
  if (s.eff < 0.6) return s;
  s.patch_rows = s.pbw * s.pbh * s.pbt;
  // weights resident if they fit next to two patches, otherwise streamed per channel tile
  // (resident mode signals each tap's weights on one of the 2 * kPatchMaxBStages stream barriers)
  // A trailing 16 or 32 channels (conv2 temporal: 144 = 2 x 64 + 16, stem temporal: 96 = 64 + 32)
  // go into a narrow tail block instead of a zero-filled 64-channel one (single-CTA mode only).
  const int cp = g.cin_pad();
  const int tail = (!pair && cp > 64 && (cp % 64 == 16 || cp % 64 == 32)) ? cp % 64 : 0;
  const int nkb_res = tail ? cp / 64 : s.nkb;
  if (np <= kMaxBlockN && s.taps <= 2 * kPatchMaxBStages &&
      conv_patch_smem_bytes(nkb_res, s.patch_rows, s.taps, np / ncta, 0, 2, tail) <= kSmemLimit) {
    s.nkb = nkb_res;
    s.tail_ch = tail;
    s.n_tiles = 1;
    s.block_n = np;
    s.b_stages = 0;
    while (s.p_bufs < kPatchMaxBufs &&
           conv_patch_smem_bytes(s.nkb, s.patch_rows, s.taps, np / ncta, 0, s.p_bufs + 1, tail) <=
               kSmemLimit)
      ++s.p_bufs;
    s.ok = true;
    return s;
  }
  for (int nt = ceil_div(np, kMaxBlockN); nt <= 8; ++nt) {
    const int bn = round_up(ceil_div(np, nt), 16);
    const int stage = (bn / ncta) * 128;  // a CTA of a pair holds half of every weight tile
    for (int pbufs = 3; pbufs >= 2; --pbufs) {
      const int fixed =
          conv_patch_smem_bytes(s.nkb, s.patch_rows, s.taps, bn / ncta, 1, pbufs) - stage;
      int stages = (kSmemLimit - fixed) / stage;
      if (stages > kPatchMaxBStages) stages = kPatchMaxBStages;
      if (stages >= 4 && bn >= 64) {
        s.n_tiles = ceil_div(np, bn);
        s.block_n = bn;
        s.b_stages = stages;
        s.p_bufs = pbufs;
        s.ok = true;
        return s;
      }
    }
  }
  return s;
}
}  // namespace

void stem_fused_add_temporal_bias(const float* bias, int cout, std::vector<uint16_t>* packed) {
  // row n, centre tap (dt = 1), channel 95 = the constant-one channel of the intermediate
  for (int n = 0; n < cout && n < kStemOut; ++n)
    (*packed)[(size_t)n * 3 * kStemMid + kStemMid + (kStemMid - 1)] = f32_to_bf16_bits(bias[n]);
}

void pack_stem_fused_weights(const float* w, const float* scale, const float* shift, int cout,
                             std::vector<uint16_t>* packed) {
  packed->assign((size_t)kStemMid * 256, 0);
  for (int n = 0; n < cout && n < kStemMid; ++n) {
    const float sc = scale ? scale[n] : 1.f;
    // centre tap (ah, aw) = (2, 2), slot 3 of pixel (0, 0): multiplies the constant one
    if (shift) (*packed)[(size_t)n * 256 + (2 * 4 + 2) * 16 + 3] = f32_to_bf16_bits(shift[n]);
    for (int ah = 0; ah < 4; ++ah)
      for (int aw = 0; aw < 4; ++aw)
        for (int py = 0; py < 2; ++py)
          for (int px = 0; px < 2; ++px) {
            // input row 2 ho + dh - 3 = 2 (ho + ah - 2) + py
            const int dh = 2 * (ah - 2) + py + 3, dw = 2 * (aw - 2) + px + 3;
            if (dh < 0 || dw < 0) continue;
            for (int c = 0; c < 3; ++c)
              (*packed)[(size_t)n * 256 + (ah * 4 + aw) * 16 + (py * 2 + px) * 4 + c] =
                  f32_to_bf16_bits(w[(((size_t)n * 3 + c) * 7 + dh) * 7 + dw] * sc);
          }
  }
}

bool build_stem_fused_params(int B, const __nv_bfloat16* s2d, const __nv_bfloat16* w_spatial,
                             const __nv_bfloat16* w_temporal, __nv_bfloat16* y,
                             StemFusedParams* out, std::string* err) {
  StemFusedParams& p = *out;
  memset(&p, 0, sizeof p);
  p.y = y;
  p.B = B;
  {
    // dimension order (c, w, t, h, b): the box then lies in shared memory as rows h * 128 + t * 16 + w
    const uint64_t row = (uint64_t)kStemS2dC * 2, plane = (uint64_t)kStemHW * kStemHW * row;
    const uint64_t dims[5] = {(uint64_t)kStemS2dC, (uint64_t)kStemHW, (uint64_t)kStemT,
                              (uint64_t)kStemHW, (uint64_t)B};
    const uint64_t strides[4] = {row, plane, (uint64_t)kStemHW * row, (uint64_t)kStemT * plane};
    const uint32_t box[5] = {(uint32_t)kStemS2dC, 16u, (uint32_t)kStemT, 4u, 1u};
    if (!encode_bf16(&p.in_map, s2d, 5, dims, strides, box, 32, err)) return false;
  }
  {
    const uint64_t dims[2] = {256, (uint64_t)kStemMid};
    const uint64_t strides[1] = {512};
    const uint32_t box[2] = {16u, (uint32_t)kStemMid};
    if (!encode_bf16(&p.wsp_map, w_spatial, 2, dims, strides, box, 32, err)) return false;
  }
  {
    const uint64_t dims[2] = {3 * (uint64_t)kStemMid, (uint64_t)kStemOut};
    const uint64_t strides[1] = {3 * (uint64_t)kStemMid * 2};
    const uint32_t box[2] = {64u, (uint32_t)kStemOut};
    if (!encode_bf16(&p.wtp_map, w_temporal, 2, dims, strides, box, 128, err)) return false;
    const uint32_t tbox[2] = {32u, (uint32_t)kStemOut};
    if (!encode_bf16(&p.wtp_tail_map, w_temporal, 2, dims, strides, tbox, 64, err)) return false;
  }
  return true;
}

bool conv_patch_applicable(const ConvGeom& g, int pair) { return patch_shape(g, pair).ok; }
bool conv_patch_streams(const ConvGeom& g, int pair) {
  const PatchShape s = patch_shape(g, pair);
  return s.ok && s.b_stages > 0;
}

bool build_conv_patch_params(const ConvGeom& g, int B, const __nv_bfloat16* x,
                             const __nv_bfloat16* w_packed, __nv_bfloat16* y, const float* shift,
                             const __nv_bfloat16* residual, bool relu, ConvPatchParams* out,
                             std::string* err, int pair) {
  const PatchShape s = patch_shape(g, pair);
  if (!s.ok) {
    if (err) *err = "conv_patch: geometry not supported";
    return false;
  }
  ConvPatchParams& p = *out;
  memset(&p, 0, sizeof p);
  const int cp = g.cin_pad(), np = g.cout_pad();
  p.shift = shift;
  p.residual = residual;
  p.y = y;
  p.nkb = s.nkb;
  p.tail_ch = s.tail_ch;
  p.cin_pad = cp;
  p.patch_rows = s.patch_rows;
  p.num_taps = s.taps;
  p.n = np;
  p.n_tiles = s.n_tiles;
  p.block_n = s.block_n;
  p.b_stages = s.b_stages;
  p.p_bufs = s.p_bufs;
  p.pair = pair ? 1 : 0;
  p.tw = s.tw; p.th = s.th; p.tt = s.tt;
  p.lpw = s.lpw; p.lph = s.lph;
  p.tiles_w = ceil_div(g.Wo, s.tw);
  p.tiles_h = ceil_div(g.Ho, s.th);
  p.tiles_t = ceil_div(g.To, s.tt);
  p.tiles_b = B;
  p.aw = g.stem_unfolded ? 0 : (s.spatial ? -1 : 0);
  p.ah = g.stem_unfolded ? -2 : (s.spatial ? -1 : 0);
  p.at = s.spatial ? 0 : -1;
  p.Wo = g.Wo; p.Ho = g.Ho; p.To = g.To; p.B = B;
  p.cout_pad = np;
  p.relu = relu ? 1 : 0;
  for (int tap = 0; tap < s.taps; ++tap)
    p.tap_shift[tap] = (int16_t)(g.stem_unfolded ? tap * s.pbw
                                 : s.spatial    ? (tap / 3) * s.pbw + (tap % 3)
                                                : tap * s.pbw * s.pbh);
  {
    // extent of the input the patches are cut from (the unfolded stem: Wo columns, Hi/2 row pairs)
    const uint64_t in_w = g.stem_unfolded ? (uint64_t)g.Wo : (uint64_t)g.Wi;
    const uint64_t in_h = g.stem_unfolded ? (uint64_t)(g.Hi / 2) : (uint64_t)g.Hi;
    const uint64_t dims[5] = {(uint64_t)cp, in_w, in_h, (uint64_t)g.Ti, (uint64_t)B};
    const uint64_t strides[4] = {(uint64_t)cp * 2, in_w * cp * 2, in_h * in_w * cp * 2,
                                 (uint64_t)g.Ti * in_h * in_w * cp * 2};
    const uint32_t box[5] = {64u, (uint32_t)s.pbw, (uint32_t)s.pbh, (uint32_t)s.pbt, 1u};
    if (!encode_bf16(&p.a_map, x, 5, dims, strides, box, 128, err)) return false;
    p.a_map_tail = p.a_map;
    if (s.tail_ch) {
      const uint32_t tbox[5] = {(uint32_t)s.tail_ch, (uint32_t)s.pbw, (uint32_t)s.pbh,
                                (uint32_t)s.pbt, 1u};
      if (!encode_bf16(&p.a_map_tail, x, 5, dims, strides, tbox, s.tail_ch * 2, err)) return false;
    }
  }
  {
    const uint64_t dims[2] = {(uint64_t)g.k_total(), (uint64_t)np};
    const uint64_t strides[1] = {(uint64_t)g.k_total() * 2};
    const uint32_t box[2] = {64u, (uint32_t)(s.block_n / (pair ? 2 : 1))};
    if (!encode_bf16(&p.b_map, w_packed, 2, dims, strides, box, 128, err)) return false;
    p.b_map_tail = p.b_map;
    if (s.tail_ch) {
      const uint32_t tbox[2] = {(uint32_t)s.tail_ch, (uint32_t)s.block_n};
      if (!encode_bf16(&p.b_map_tail, w_packed, 2, dims, strides, tbox, s.tail_ch * 2, err))
        return false;
    }
  }
  // Residual through shared memory (resident single-CTA mode, 64 or 128 output channels): two
  // TMA-fed tiles of 128 rows, laid out like the accumulator rows.  RNB_B200_RES_SMEM=0 disables.
  p.r_map = p.a_map;
  static const bool allow_res_smem = [] {
    const char* e = getenv("RNB_B200_RES_SMEM");
    return !(e && e[0] == '0');
  }();
  if (allow_res_smem && residual && !pair && s.b_stages == 0 && (np == 64 || np == 128) &&
      s.taps <= 2 * kPatchMaxBStages - 4) {
    int bufs = s.p_bufs;
    while (bufs > 2 && conv_patch_smem_bytes(s.nkb, s.patch_rows, s.taps, np, 0, bufs, s.tail_ch,
                                             np / 64) > kSmemLimit)
      --bufs;
    if (conv_patch_smem_bytes(s.nkb, s.patch_rows, s.taps, np, 0, bufs, s.tail_ch, np / 64) <=
        kSmemLimit) {
      const int rt = 128 >> (s.lpw + s.lph);
      const uint64_t dims[5] = {(uint64_t)np, (uint64_t)g.Wo, (uint64_t)g.Ho, (uint64_t)g.To,
                                (uint64_t)B};
      const uint64_t strides[4] = {(uint64_t)np * 2, (uint64_t)g.Wo * np * 2,
                                   (uint64_t)g.Ho * g.Wo * np * 2,
                                   (uint64_t)g.To * g.Ho * g.Wo * np * 2};
      const uint32_t box[5] = {64u, 1u << s.lpw, 1u << s.lph, (uint32_t)rt, 1u};
      if (!encode_bf16(&p.r_map, residual, 5, dims, strides, box, 128, err)) return false;
      p.res_smem = 1;
      p.p_bufs = bufs;
    }
  }
  return true;
}

// ---- temporal convolution with the taps along N (conv_tsum.cuh)
bool conv_tsum_applicable(const ConvGeom& g) {
  const int cp = g.cin_pad(), tail = cp % 64;
  return !g.stem_unfolded && g.kt == 3 && g.kh == 1 && g.kw == 1 && g.st == 1 && g.sh == 1 &&
         g.sw == 1 && g.pt == 1 && g.ph == 0 && g.pw == 0 && g.Ti == kTsumT && g.To == kTsumT &&
         g.cout_pad() == kTsumCout && g.extra_cin == 0 && g.ncat == 0 &&
         (g.in_pitch == 0 || g.in_pitch == cp) && cp <= kTsumMaxKb * 64 &&
         (tail == 0 || tail == 16 || tail == 32 || cp + (64 - tail) <= kTsumMaxKb * 64);
}

bool build_conv_tsum_params(const ConvGeom& g, int B, const __nv_bfloat16* x,
                            const __nv_bfloat16* w_packed, __nv_bfloat16* y, const float* shift,
                            const __nv_bfloat16* residual, bool relu, ConvTsumParams* out,
                            std::string* err) {
  if (!conv_tsum_applicable(g)) {
    if (err) *err = "conv_tsum: geometry not supported";
    return false;
  }
  ConvTsumParams& p = *out;
  memset(&p, 0, sizeof p);
  const int cp = g.cin_pad();
  const int tail = cp % 64;
  // 48 trailing channels: a zero-filled 64-channel block (the tensor map's bound) instead of a tail
  p.tail_ch = (tail == 16 || tail == 32) ? tail : 0;
  p.nkb = p.tail_ch ? cp / 64 : ceil_div(cp, 64);
  p.cin_pad = cp;
  p.HW = g.Ho * g.Wo;
  p.tiles_hw = ceil_div(p.HW, kTsumPos);
  p.B = B;
  p.shift = shift;
  p.residual = residual;
  p.y = y;
  p.relu = relu ? 1 : 0;
  p.r_bufs = residual ? 3 : 0;
  p.p_bufs = 6;
  while (p.p_bufs > 2 && conv_tsum_smem_bytes(p.nkb, p.tail_ch, p.p_bufs, p.r_bufs) > kSmemLimit)
    --p.p_bufs;
  if (conv_tsum_smem_bytes(p.nkb, p.tail_ch, p.p_bufs, p.r_bufs) > kSmemLimit) {
    if (err) *err = "conv_tsum: does not fit in shared memory";
    return false;
  }
  {
    const uint64_t hw = (uint64_t)p.HW, c2 = (uint64_t)cp * 2;
    const uint64_t dims[5] = {(uint64_t)cp, (uint64_t)kTsumT, hw, (uint64_t)B, 1};
    const uint64_t strides[4] = {hw * c2, c2, (uint64_t)kTsumT * hw * c2,
                                 (uint64_t)B * kTsumT * hw * c2};
    const uint32_t box[5] = {64u, (uint32_t)kTsumT, (uint32_t)kTsumPos, 1u, 1u};
    if (!encode_bf16(&p.a_map, x, 5, dims, strides, box, 128, err)) return false;
    p.r_map = p.a_map;
    if (residual) {
      const uint64_t o2 = (uint64_t)kTsumCout * 2;
      const uint64_t rdims[5] = {(uint64_t)kTsumCout, (uint64_t)kTsumT, hw, (uint64_t)B, 1};
      const uint64_t rstrides[4] = {hw * o2, o2, (uint64_t)kTsumT * hw * o2,
                                    (uint64_t)B * kTsumT * hw * o2};
      if (!encode_bf16(&p.r_map, residual, 5, rdims, rstrides, box, 128, err)) return false;
    }
    p.a_map_tail = p.a_map;
    if (p.tail_ch) {
      const uint32_t tbox[5] = {(uint32_t)p.tail_ch, (uint32_t)kTsumT, (uint32_t)kTsumPos, 1u, 1u};
      if (!encode_bf16(&p.a_map_tail, x, 5, dims, strides, tbox, p.tail_ch * 2, err)) return false;
    }
  }
  {
    const uint64_t dims[2] = {(uint64_t)g.k_total(), (uint64_t)kTsumCout};
    const uint64_t strides[1] = {(uint64_t)g.k_total() * 2};
    const uint32_t box[2] = {64u, (uint32_t)kTsumCout};
    if (!encode_bf16(&p.b_map, w_packed, 2, dims, strides, box, 128, err)) return false;
    p.b_map_tail = p.b_map;
    if (p.tail_ch) {
      const uint32_t tbox[2] = {(uint32_t)p.tail_ch, (uint32_t)kTsumCout};
      if (!encode_bf16(&p.b_map_tail, w_packed, 2, dims, strides, tbox, p.tail_ch * 2, err))
        return false;
    }
  }
  return true;
}

}  // namespace rnb
