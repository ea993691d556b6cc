// R(2+1)D layer-range model on top of the convolution kernels.
//
// Structure follows the reference's R2Plus1DLayerNet / R2Plus1DLayerWrapper
// (models/r2p1d/network.py:9-60) and the un-vendored irhum/R2Plus1D-PyTorch modules they
// instantiate (see oracle/r2p1d_ref.py for what is pinned and what is assumed).  Every
// SpatioTemporalConv becomes two convolution launches (1 x k x k, then k x 1 x 1) whose epilogues
// carry the folded BatchNorm / bias, the ReLU and, for the second convolution of a residual block,
// the shortcut add.
#include "model.cuh"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/rnb_b200.h"

namespace rnb {

namespace {
constexpr float kBnEps = 1e-5f;  // torch.nn.BatchNorm3d default

std::string shape_str(const std::vector<int64_t>& s) {
  std::string r = "(";
  for (size_t i = 0; i < s.size(); ++i) r += (i ? ", " : "") + std::to_string(s[i]);
  return r + ")";
}

bool cuda_ok(cudaError_t e, const char* what, std::string* err) {
  if (e == cudaSuccess) return true;
  if (err) *err = std::string(what) + ": " + cudaGetErrorString(e);
  return false;
}

int mid_channels(int cin, int cout, int kt, int kh, int kw) {
  return (int)floor((double)(kt * kh * kw) * cin * cout / (double)(kh * kw * cin + kt * cout));
}
}  // namespace

R2P1DModel::R2P1DModel(std::vector<int> layer_sizes, int start_idx, int end_idx, int num_classes,
                       int dtype, int max_batch)
    : layer_sizes_(std::move(layer_sizes)),
      start_idx_(start_idx),
      end_idx_(end_idx),
      num_classes_(num_classes),
      dtype_(dtype),
      max_batch_(max_batch) {
  static const Dims kIn[6] = {{}, {3, 8, 112, 112}, {64, 8, 56, 56}, {64, 8, 56, 56},
                              {128, 4, 28, 28}, {256, 2, 14, 14}};
  static const Dims kOut[6] = {{}, {64, 8, 56, 56}, {64, 8, 56, 56}, {128, 4, 28, 28},
                               {256, 2, 14, 14}, {512, 1, 7, 7}};
  if (const char* e = getenv("RNB_B200_FUSE_SHORTCUT")) fuse_shortcut_ = e[0] != '0';
  stem_fused_ = dtype == RNB_DTYPE_BF16 && start_idx == 1;
  if (const char* e = getenv("RNB_B200_STEM")) stem_fused_ = stem_fused_ && strcmp(e, "unfolded") != 0;
  in_dims_ = kIn[start_idx];
  out_dims_ = kOut[end_idx];
}

R2P1DModel::~R2P1DModel() { release(); }

void R2P1DModel::release() {
  if (finalized_) {
    // (the L2 carve-out is a device-wide limit: the last handle on the device puts it back)
    int cur = 0;
    if (cudaGetDevice(&cur) == cudaSuccess) {
      if (cur != device_) cudaSetDevice(device_);
      l2_persist_release();
      if (cur != device_) cudaSetDevice(cur);
    }
    finalized_ = false;
  }
  for (ConvOp& op : ops_) {
    if (op.d_w) cudaFree(op.d_w);
    if (op.d_scale) cudaFree(op.d_scale);
    if (op.d_shift) cudaFree(op.d_shift);
    op.d_w = nullptr;
    op.d_scale = op.d_shift = nullptr;
  }
  for (void*& s : slots_) {
    if (s) cudaFree(s);
    s = nullptr;
  }
  if (d_classify_scratch_) cudaFree(d_classify_scratch_);
  d_classify_scratch_ = nullptr;
  if (d_dual_flags_) cudaFree(d_dual_flags_);
  d_dual_flags_ = nullptr;
  if (d_stem_w_) cudaFree(d_stem_w_);
  d_stem_w_ = nullptr;
  if (d_fc_w_) cudaFree(d_fc_w_);
  if (d_fc_b_) cudaFree(d_fc_b_);
  d_fc_w_ = d_fc_b_ = nullptr;
}

bool R2P1DModel::load_param(const char* key, const float* data, const int64_t* shape, int ndim,
                            std::string* err) {
  if (finalized_) {
    if (err) *err = "load_param after finalize";
    return false;
  }
  HostTensor t;
  int64_t n = 1;
  for (int i = 0; i < ndim; ++i) {
    t.shape.push_back(shape[i]);
    n *= shape[i];
  }
  t.data.assign(data, data + n);
  params_[key] = std::move(t);
  return true;
}

const HostTensor* R2P1DModel::get(const std::string& key, std::vector<int64_t> shape,
                                  std::string* err) const {
  auto it = params_.find(key);
  if (it == params_.end()) {
    if (err) *err = "missing state_dict entry '" + key + "'";
    return nullptr;
  }
  if (it->second.shape != shape) {
    if (err)
      *err = "state_dict entry '" + key + "' has shape " + shape_str(it->second.shape) +
             ", expected " + shape_str(shape);
    return nullptr;
  }
  return &it->second;
}

// y = gamma * (x + bias - mean) / sqrt(var + eps) + beta  ->  scale, shift applied to the raw conv
bool R2P1DModel::fold_bn(const std::string& bn, const std::vector<float>& bias,
                         std::vector<float>* scale, std::vector<float>* shift,
                         std::string* err) const {
  const int64_t n = (int64_t)bias.size();
  scale->assign(n, 1.f);
  *shift = bias;
  if (bn.empty()) return true;
  const HostTensor* gamma = get(bn + ".weight", {n}, err);
  const HostTensor* beta = gamma ? get(bn + ".bias", {n}, err) : nullptr;
  const HostTensor* mean = beta ? get(bn + ".running_mean", {n}, err) : nullptr;
  const HostTensor* var = mean ? get(bn + ".running_var", {n}, err) : nullptr;
  if (!var) return false;
  for (int64_t i = 0; i < n; ++i) {
    const float s = gamma->data[i] / sqrtf(var->data[i] + kBnEps);
    (*scale)[i] = s;
    (*shift)[i] = beta->data[i] + (bias[i] - mean->data[i]) * s;
  }
  return true;
}

bool R2P1DModel::add_st_conv(const std::string& prefix, int in_slot, int out_slot, Dims in, int cout,
                             int kt, int kh, int kw, int st, int sh, int sw, int pt, int ph, int pw,
                             const std::string& post_bn, bool relu, int res_slot, bool stem,
                             Dims* out, std::string* err) {
  const int mid = mid_channels(in.C, cout, kt, kh, kw);
  // ---- spatial 1 x kh x kw, then the module's own BatchNorm + ReLU
  ConvOp sp;
  sp.name = prefix + ".spatial_conv";
  sp.g.Ti = in.T; sp.g.Hi = in.H; sp.g.Wi = in.W; sp.g.cin = in.C; sp.g.cout = mid;
  sp.g.kt = 1; sp.g.kh = kh; sp.g.kw = kw;
  sp.g.st = 1; sp.g.sh = sh; sp.g.sw = sw;
  sp.g.pt = 0; sp.g.ph = ph; sp.g.pw = pw;
  sp.g.To = in.T;
  sp.g.Ho = (in.H + 2 * ph - kh) / sh + 1;
  sp.g.Wo = (in.W + 2 * pw - kw) / sw + 1;
  sp.g.stem_unfolded = stem && dtype_ == RNB_DTYPE_BF16;
  sp.in_slot = in_slot;
  sp.out_slot = SLOT_M;
  sp.relu = true;
  const HostTensor* w = get(sp.name + ".weight", {mid, in.C, 1, kh, kw}, err);
  const HostTensor* b = w ? get(sp.name + ".bias", {mid}, err) : nullptr;
  if (!b) return false;
  sp.w = w->data;
  if (!fold_bn(prefix + ".bn", b->data, &sp.scale, &sp.shift, err)) return false;
  // ---- temporal kt x 1 x 1, then whatever BatchNorm / ReLU / residual the caller attaches
  ConvOp tp;
  tp.name = prefix + ".temporal_conv";
  tp.g.Ti = sp.g.To; tp.g.Hi = sp.g.Ho; tp.g.Wi = sp.g.Wo; tp.g.cin = mid; tp.g.cout = cout;
  tp.g.kt = kt; tp.g.kh = 1; tp.g.kw = 1;
  tp.g.st = st; tp.g.sh = 1; tp.g.sw = 1;
  tp.g.pt = pt; tp.g.ph = 0; tp.g.pw = 0;
  tp.g.To = (tp.g.Ti + 2 * pt - kt) / st + 1;
  tp.g.Ho = tp.g.Hi;
  tp.g.Wo = tp.g.Wi;
  tp.in_slot = SLOT_M;
  tp.out_slot = out_slot;
  tp.res_slot = res_slot;
  tp.relu = relu;
  w = get(tp.name + ".weight", {cout, mid, kt, 1, 1}, err);
  b = w ? get(tp.name + ".bias", {cout}, err) : nullptr;
  if (!b) return false;
  tp.w = w->data;
  if (!fold_bn(post_bn, b->data, &tp.scale, &tp.shift, err)) return false;
  *out = {cout, tp.g.To, tp.g.Ho, tp.g.Wo};
  ops_.push_back(std::move(sp));
  ops_.push_back(std::move(tp));
  return true;
}

bool R2P1DModel::build_ops(std::string* err) {
  ops_.clear();
  int cur = SLOT_X, nxt = SLOT_Y;
  Dims d = in_dims_;
  static const int kStageOut[6] = {0, 64, 64, 128, 256, 512};
  for (int layer = start_idx_, pos = 0; layer <= end_idx_; ++layer, ++pos) {
    const std::string lp = "res2plus1d.conv" + std::to_string(layer);
    if (layer == 1) {
      // SpatioTemporalConv(3, 64, [3,7,7], stride=[1,2,2], padding=[1,3,3]); network.py:22.
      // No BatchNorm / ReLU follows the stem (assumed from the public submodule).
      if (!add_st_conv(lp, SLOT_U, cur, d, 64, 3, 7, 7, 1, 2, 2, 1, 3, 3, "", false, -1, true, &d,
                       err))
        return false;
      continue;
    }
    const int n_blocks = layer_sizes_.empty() ? 2 : layer_sizes_[pos];
    if (n_blocks < 1) {
      if (err) *err = "layer_sizes entries of residual stages must be >= 1";
      return false;
    }
    const bool down = layer >= 3;  // network.py:28,31,34
    const int cout = kStageOut[layer];
    for (int blk = 0; blk < n_blocks; ++blk) {
      const std::string bp =
          blk == 0 ? lp + ".block1" : lp + ".blocks." + std::to_string(blk - 1);
      const bool ds = down && blk == 0;
      int res = cur;
      Dims dres = d, d1 = d, d2 = d;
      if (ds) {
        if (!add_st_conv(bp + ".downsampleconv", cur, SLOT_S, d, cout, 1, 1, 1, 2, 2, 2, 0, 0, 0,
                         bp + ".downsamplebn", false, -1, false, &dres, err))
          return false;
        res = SLOT_S;
      }
      const int s = ds ? 2 : 1;
      if (!add_st_conv(bp + ".conv1", cur, SLOT_R, d, cout, 3, 3, 3, s, s, s, 1, 1, 1, bp + ".bn1",
                       true, -1, false, &d1, err))
        return false;
      if (!add_st_conv(bp + ".conv2", SLOT_R, nxt, d1, cout, 3, 3, 3, 1, 1, 1, 1, 1, 1, bp + ".bn2",
                       true, res, false, &d2, err))
        return false;
      if (ds && dtype_ == RNB_DTYPE_BF16 && fuse_shortcut_) {
        // relu(bn2(conv2(x)) + bn_ds(conv_ds(x))): the shortcut's 1x1x1 temporal convolution is
        // linear in its input, so it becomes one more K slice of conv2's temporal GEMM (its own
        // BatchNorm scale folded into its weight rows, the two shifts added) instead of a launch
        // of its own plus a residual read.  Its input, the shortcut's spatial output, moves to a
        // slot of its own because slot M is reused before conv2's temporal convolution runs.
        const size_t n = ops_.size();  // [.. ds.sp, ds.tp, c1.sp, c1.tp, c2.sp, c2.tp]
        ConvOp& ds_sp = ops_[n - 6];
        ConvOp& ds_tp = ops_[n - 5];
        ConvOp& c1_sp = ops_[n - 4];
        ConvOp& c1_tp = ops_[n - 3];
        ConvOp& c2_tp = ops_[n - 1];
        // The shortcut's 1x1x1 SPATIAL convolution (stride 1,2,2) reads exactly the input positions
        // the centre tap of conv1's 1x3x3 (stride 1,2,2) reads, and both are followed by their own
        // BatchNorm + ReLU: it becomes extra output channels of conv1's spatial GEMM (zero weights
        // on the other eight taps).  The widened output lives in slot D; conv1's temporal
        // convolution reads its first pad16(mid) channels, the fused tap below the rest.
        const int mid1_pad = chan_pad(c1_sp.g.cout);
        const int ds_pad = chan_pad(ds_sp.g.cout);
        const int wide = mid1_pad + ds_pad;
        c1_sp.g.ncat = ds_sp.g.cout;
        c1_sp.w_ncat = ds_sp.w;
        c1_sp.scale_ncat = ds_sp.scale;
        c1_sp.scale.resize(wide, 0.f);
        c1_sp.shift.resize(wide, 0.f);
        for (int i = 0; i < ds_sp.g.cout; ++i) {
          c1_sp.scale[mid1_pad + i] = ds_sp.scale[i];
          c1_sp.shift[mid1_pad + i] = ds_sp.shift[i];
        }
        c1_sp.out_slot = SLOT_D;
        c1_sp.name += "+shortcut";
        c1_tp.in_slot = SLOT_D;
        c1_tp.g.in_pitch = wide;
        c2_tp.g.extra_cin = ds_tp.g.cin;
        c2_tp.g.extra_st = ds_tp.g.st;
        c2_tp.g.extra_Ti = ds_tp.g.Ti;
        c2_tp.g.extra_pitch = wide;
        c2_tp.g.extra_coff = mid1_pad;
        c2_tp.extra_slot = SLOT_D;
        c2_tp.w_extra = ds_tp.w;
        c2_tp.scale_extra = ds_tp.scale;
        for (size_t i = 0; i < c2_tp.shift.size(); ++i) c2_tp.shift[i] += ds_tp.shift[i];
        c2_tp.res_slot = -1;
        c2_tp.name += "+shortcut";
        ops_.erase(ops_.begin() + (n - 6), ops_.begin() + (n - 4));
      }
      d = d2;
      std::swap(cur, nxt);
    }
  }
  final_slot_ = cur;
  if (d.C != out_dims_.C || d.T != out_dims_.T || d.H != out_dims_.H || d.W != out_dims_.W) {
    if (err) *err = "internal: activation shape does not match the reference's shape table";
    return false;
  }
  return true;
}

int R2P1DModel::finalize(std::string* err) {
  if (finalized_) return RNB_OK;
  if (!cuda_ok(cudaGetDevice(&device_), "cudaGetDevice", err)) return RNB_ECUDA;
  if (!cuda_ok(cudaDeviceGetAttribute(&num_sms_, cudaDevAttrMultiProcessorCount, device_),
               "cudaDeviceGetAttribute", err))
    return RNB_ECUDA;
  if (!build_ops(err)) return RNB_EPARAM;

  // ---- arena: every slot sized for its largest tenant at max_batch
  memset(slot_elems_per_clip_, 0, sizeof slot_elems_per_clip_);
  auto need = [&](int slot, size_t elems) {
    if (elems > slot_elems_per_clip_[slot]) slot_elems_per_clip_[slot] = elems;
  };
  if (start_idx_ == 1) {
    need(SLOT_U, dtype_ == RNB_DTYPE_BF16 ? (size_t)kClipT * kClipH * (kClipW / 2) * kStemUnfold
                                          : (size_t)kClipT * kClipH * kClipW * kClipC);
  } else {
    need(SLOT_X, (size_t)in_dims_.T * in_dims_.H * in_dims_.W * chan_pad(in_dims_.C));
  }
  for (const ConvOp& op : ops_)
    need(op.out_slot, (size_t)op.g.To * op.g.Ho * op.g.Wo *
                          (dtype_ == RNB_DTYPE_BF16 ? op.g.cout_pad() : op.g.cout));
  for (int s = 0; s < NUM_SLOTS; ++s) {
    if (!slot_elems_per_clip_[s]) continue;
    const size_t bytes = slot_elems_per_clip_[s] * (size_t)max_batch_ * elem_size();
    if (!cuda_ok(cudaMalloc(&slots_[s], bytes), "cudaMalloc(activation arena)", err)) {
      release();
      return RNB_ECUDA;
    }
  }

  // ---- parameters
  for (ConvOp& op : ops_) {
    const int np = dtype_ == RNB_DTYPE_BF16 ? op.g.cout_pad() : op.g.cout;
    std::vector<float> sc(np, 0.f), sh(np, 0.f);
    memcpy(sc.data(), op.scale.data(), op.scale.size() * 4);
    memcpy(sh.data(), op.shift.data(), op.shift.size() * 4);
    bool ok = cuda_ok(cudaMalloc(&op.d_scale, np * 4), "cudaMalloc", err) &&
              cuda_ok(cudaMalloc(&op.d_shift, np * 4), "cudaMalloc", err) &&
              cuda_ok(cudaMemcpy(op.d_scale, sc.data(), np * 4, cudaMemcpyHostToDevice), "cudaMemcpy", err) &&
              cuda_ok(cudaMemcpy(op.d_shift, sh.data(), np * 4, cudaMemcpyHostToDevice), "cudaMemcpy", err);
    if (ok && stem_fused_ && &op == &ops_[0]) {
      std::vector<uint16_t> fused;
      pack_stem_fused_weights(op.w.data(), op.scale.data(), op.shift.data(), op.g.cout, &fused);
      ok = cuda_ok(cudaMalloc(&d_stem_w_, fused.size() * 2), "cudaMalloc(stem weights)", err) &&
           cuda_ok(cudaMemcpy(d_stem_w_, fused.data(), fused.size() * 2, cudaMemcpyHostToDevice),
                   "cudaMemcpy(stem weights)", err);
    }
    if (ok && dtype_ == RNB_DTYPE_BF16) {
      std::vector<uint16_t> packed;
      pack_conv_weights_tc(op.g, op.w.data(), op.scale.data(), &packed,
                           op.w_extra.empty() ? nullptr : op.w_extra.data(),
                           op.scale_extra.empty() ? nullptr : op.scale_extra.data(),
                           op.w_ncat.empty() ? nullptr : op.w_ncat.data(),
                           op.scale_ncat.empty() ? nullptr : op.scale_ncat.data());
      // fused stem: the temporal bias rides in the GEMM (conv_stem.cuh)
      if (stem_fused_ && &op == &ops_[1])
        stem_fused_add_temporal_bias(op.shift.data(), op.g.cout, &packed);
      ok = cuda_ok(cudaMalloc(&op.d_w, packed.size() * 2), "cudaMalloc(weights)", err) &&
           cuda_ok(cudaMemcpy(op.d_w, packed.data(), packed.size() * 2, cudaMemcpyHostToDevice),
                   "cudaMemcpy(weights)", err);
    } else if (ok) {
      std::vector<float> packed;
      pack_conv_weights_ref(op.g, op.w.data(), &packed);
      ok = cuda_ok(cudaMalloc(&op.d_w, packed.size() * 4), "cudaMalloc(weights)", err) &&
           cuda_ok(cudaMemcpy(op.d_w, packed.data(), packed.size() * 4, cudaMemcpyHostToDevice),
                   "cudaMemcpy(weights)", err);
    }
    if (!ok) {
      release();
      return RNB_ECUDA;
    }
    op.w.clear();
    op.w.shrink_to_fit();
  }
  if (end_idx_ == 5) {
    const HostTensor* w = get("linear.weight", {num_classes_, 512}, err);
    const HostTensor* b = w ? get("linear.bias", {num_classes_}, err) : nullptr;
    if (!b) {
      release();
      return RNB_EPARAM;
    }
    bool ok = cuda_ok(cudaMalloc(&d_classify_scratch_,
                                 pool_fc_classify_scratch_bytes(max_batch_, num_classes_)),
                      "cudaMalloc(classifier scratch)", err) &&
              cuda_ok(cudaMalloc(&d_fc_w_, w->data.size() * 4), "cudaMalloc", err) &&
              cuda_ok(cudaMalloc(&d_fc_b_, b->data.size() * 4), "cudaMalloc", err) &&
              cuda_ok(cudaMemcpy(d_fc_w_, w->data.data(), w->data.size() * 4, cudaMemcpyHostToDevice), "cudaMemcpy", err) &&
              cuda_ok(cudaMemcpy(d_fc_b_, b->data.data(), b->data.size() * 4, cudaMemcpyHostToDevice), "cudaMemcpy", err);
    if (!ok) {
      release();
      return RNB_ECUDA;
    }
  }
  if (dtype_ == RNB_DTYPE_BF16) {
    const size_t fb = (size_t)(max_batch_ + 2) * sizeof(unsigned long long);
    if (!cuda_ok(cudaMalloc(&d_dual_flags_, fb), "cudaMalloc(counters)", err) ||
        !cuda_ok(cudaMemset(d_dual_flags_, 0, fb), "cudaMemset(counters)", err)) {
      release();
      return RNB_ECUDA;
    }
  }
  params_.clear();
  finalized_ = true;
  l2_persist_acquire();
  return RNB_OK;
}

void R2P1DModel::io_shape(int which, int64_t shape4[4]) const {
  if (which == 1 && end_idx_ == 5) {
    shape4[0] = num_classes_;
    shape4[1] = shape4[2] = shape4[3] = 1;
    return;
  }
  const Dims& d = which == 0 ? in_dims_ : out_dims_;
  shape4[0] = d.C;
  shape4[1] = d.T;
  shape4[2] = d.H;
  shape4[3] = d.W;
}

double R2P1DModel::flops(int batch) const {
  double f = 0;
  for (const ConvOp& op : ops_) f += op.g.flops(batch);
  if (end_idx_ == 5) f += 2.0 * batch * 512 * num_classes_;
  return f;
}

bool R2P1DModel::op_info(int op, std::string* name, double* flops_per_clip,
                         double* act_bytes_per_clip, double* weight_bytes, int* kernel_kind) const {
  if (op < 0 || op >= (int)ops_.size()) return false;
  const ConvOp& o = ops_[op];
  const double es = (double)elem_size();
  double act;
  if (dtype_ == RNB_DTYPE_BF16) {
    act = (double)o.g.in_elems(1) + (double)o.g.out_elems(1) * (o.res_slot >= 0 ? 2 : 1) +
          (double)o.g.To * o.g.Ho * o.g.Wo * o.g.extra_cin_pad();
  } else {
    act = (double)o.g.Ti * o.g.Hi * o.g.Wi * o.g.cin +
          (double)o.g.To * o.g.Ho * o.g.Wo * o.g.cout * (o.res_slot >= 0 ? 2 : 1);
  }
  if (name) *name = o.name;
  if (kernel_kind) {
    // what the launch plan of the most recent batch size chose (before any forward: what applies)
    int kind = dtype_ != RNB_DTYPE_BF16 ? 0 : (conv_patch_applicable(o.g) ? 2 : 1);
    if (stem_fused_ && op < 2) kind = 7;
    const auto it = plans_.find(last_batch_);
    if (dtype_ == RNB_DTYPE_BF16 && it != plans_.end())
      for (const Launch& l : it->second)
        if (l.op == op) {
          kind = l.patch ? 2 : (l.tc.pair ? (l.tc.tpatch ? 5 : 4) : (l.tc.ks > 1 ? 3 : 1));
          if (l.tsum) kind = 8;
          if (l.dual_n_sp > 0 || l.merged) kind = 6;
          if (l.stem_fused || (stem_fused_ && op < 2)) kind = 7;
          break;
        }
    *kernel_kind = kind;
  }
  if (flops_per_clip) *flops_per_clip = o.g.flops(1);
  if (act_bytes_per_clip) *act_bytes_per_clip = act * es;
  if (weight_bytes)
    *weight_bytes = dtype_ == RNB_DTYPE_BF16
                        ? (double)o.g.cout_pad() * o.g.k_total() * 2
                        : (double)o.g.cout * o.g.cin * o.g.kt * o.g.kh * o.g.kw * 4;
  return true;
}

int R2P1DModel::profile(const void* in, int in_fmt, int batch, void* out, int out_fmt,
                        cudaStream_t stream, float* ms, int cap, std::string* err) {
  const int n_ops = (int)ops_.size();
  if (!ms || cap < n_ops + 2) {
    if (err) *err = "profile: ms buffer too small";
    return RNB_EINVAL;
  }
  // one event per kernel; a chunked convolution's launches are summed into its entry
  std::vector<int> launch_op;
  if (dtype_ == RNB_DTYPE_BF16 && batch > 0 && batch <= max_batch_) {
    const std::vector<Launch>* plan = ensure_plan(batch, stream, err, false);
    last_batch_ = batch;
    if (!plan) return RNB_ECUDA;
    for (const Launch& l : *plan)
      if (!l.merged) launch_op.push_back(l.op);
  } else {
    for (int i = 0; i < n_ops; ++i) launch_op.push_back(i);
  }
  const int n = (int)launch_op.size() + 2;
  std::vector<cudaEvent_t> ev(n + 1, nullptr);
  for (auto& e : ev)
    if (!cuda_ok(cudaEventCreate(&e), "cudaEventCreate", err)) return RNB_ECUDA;
  int rc = forward(in, in_fmt, batch, out, out_fmt, stream, err, ev.data());
  if (rc == RNB_OK && !cuda_ok(cudaStreamSynchronize(stream), "profile sync", err)) rc = RNB_ECUDA;
  for (int i = 0; i < n_ops + 2; ++i) ms[i] = 0.f;
  if (rc == RNB_OK && batch > 0) {
    for (int i = 0; i < n; ++i) {
      float t = 0.f;
      if (!cuda_ok(cudaEventElapsedTime(&t, ev[i], ev[i + 1]), "elapsed", err)) {
        rc = RNB_ECUDA;
        break;
      }
      const int slot = i == 0 ? 0 : (i == n - 1 ? n_ops + 1 : 1 + launch_op[i - 1]);
      ms[slot] += t;
    }
  }
  for (auto& e : ev)
    if (e) cudaEventDestroy(e);
  return rc;
}

// Plans one launch.  With `like` the kernel choice (patch vs tap kernel, cluster shape) is copied
// from an already planned launch of the same layer; otherwise the candidates are timed -- which is
// fastest depends on the layer (shared-memory bandwidth vs. tile fit vs. lock-step between CTAs),
// and the reference runs with cudnn.benchmark = True for the same reason (runner.py:25).
bool R2P1DModel::plan_conv(int op_index, int first, int batch, bool chunked, cudaStream_t stream,
                           const Launch* like, Launch* out, std::string* err, bool tune) {
  const ConvOp& op = ops_[op_index];
  // a chunked spatial->temporal pair reuses the START of the intermediate slot for every chunk
  const size_t in_off = (chunked && op.in_slot == SLOT_M) ? 0 : (size_t)first * op.g.in_elems(1);
  const size_t out_off = (chunked && op.out_slot == SLOT_M) ? 0 : (size_t)first * op.g.out_elems(1);
  const __nv_bfloat16* x = (const __nv_bfloat16*)slots_[op.in_slot] + in_off;
  const __nv_bfloat16* w = (const __nv_bfloat16*)op.d_w;
  __nv_bfloat16* y = (__nv_bfloat16*)slots_[op.out_slot] + out_off;
  const __nv_bfloat16* res =
      op.res_slot >= 0 ? (const __nv_bfloat16*)slots_[op.res_slot] + out_off : nullptr;
  const __nv_bfloat16* xe =
      op.extra_slot >= 0
          ? (const __nv_bfloat16*)slots_[op.extra_slot] +
                (size_t)first * op.g.extra_Ti * op.g.Ho * op.g.Wo * op.g.extra_pitch
          : nullptr;
  out->op = op_index;
  out->first = first;
  out->batch = batch;
  std::string perr;
  auto fail = [&](const std::string& m) {
    if (err) *err = op.name + ": " + m;
    return false;
  };
  if (like && like->tsum) {
    out->patch = false;
    out->tsum = true;
    return build_conv_tsum_params(op.g, batch, x, w, y, op.d_shift, res, op.relu, &out->ts, &perr)
               ? true
               : fail(perr);
  }
  if (like) {
    out->patch = like->patch;
    const bool ok = like->patch
                        ? build_conv_patch_params(op.g, batch, x, w, y, op.d_shift, res, op.relu,
                                                  &out->pp, &perr, like->pp.pair)
                    : like->tc.tpatch
                        ? build_conv_tpatch_params(op.g, batch, x, w, y, op.d_shift, res, op.relu,
                                                   like->tc.block_n, &out->tc, &perr)
                        : build_conv_tc_params(op.g, batch, x, w, y, op.d_shift, res, op.relu,
                                               num_sms_, &out->tc, &perr, like->tc.cm, like->tc.cn,
                                               like->tc.ks, like->tc.block_n, like->tc.pair, xe);
    return ok ? true : fail(perr);
  }
  // RNB_B200_PATCH_STREAM=0 keeps streamed-weight patch launches out of the candidate set
  static const bool allow_stream = [] {
    const char* e = getenv("RNB_B200_PATCH_STREAM");
    return !(e && e[0] == '0');
  }();
  const bool patch_ok = conv_patch_applicable(op.g) && (allow_stream || !conv_patch_streams(op.g));
  const bool patch_only = patch_ok && !conv_patch_streams(op.g);
  if (patch_only) {
    out->patch = true;
    // RNB_B200_PATCH_PAIR=1 (measurements only): the CTA-pair form of the resident patch kernel on
    // the 1 x k x k layers it applies to
    static const bool want_pair = [] {
      const char* e = getenv("RNB_B200_PATCH_PAIR");
      return e && e[0] == '1';
    }();
    const int pair = want_pair && op.g.kh > 1 && conv_patch_applicable(op.g, 1) &&
                             !conv_patch_streams(op.g, 1)
                         ? 1
                         : 0;
    if (!build_conv_patch_params(op.g, batch, x, w, y, op.d_shift, res, op.relu, &out->pp, &perr,
                                 pair))
      return fail(perr);
    // 3x1x1 layers of 8-frame clips: the taps-along-N kernel (conv_tsum.cuh).  Measured equal to
    // the patch kernel within 1-2 % in the real step (profiles/r02_conv_tsum.md: both designs load
    // the SM's shared-memory data pipe about equally), and its fp32 summation order differs, so it
    // is opt-in: RNB_B200_TSUM=1 takes it wherever it applies, RNB_B200_TSUM=time lets the timing
    // decide, anything else (the default) keeps the patch kernel.
    static const int tsum_mode = [] {
      const char* e = getenv("RNB_B200_TSUM");
      return !e ? 0 : (e[0] == '1' ? 1 : (e[0] == 't' ? -1 : 0));
    }();
    ConvTsumParams ts;
    std::string terr;
    if (tsum_mode != 0 && (tune || tsum_mode == 1) && conv_tsum_applicable(op.g) &&
        build_conv_tsum_params(op.g, batch, x, w, y, op.d_shift, res, op.relu, &ts, &terr)) {
      bool use = tsum_mode == 1;
      cudaEvent_t e0, e1;
      if (!use && cudaEventCreate(&e0) == cudaSuccess) {
        if (cudaEventCreate(&e1) == cudaSuccess) {
          auto time_it = [&](auto&& launch, float* ms) -> bool {
            cudaError_t te = launch();
            if (te == cudaSuccess) te = cudaEventRecord(e0, stream);
            for (int r = 0; r < 8 && te == cudaSuccess; ++r) te = launch();
            if (te == cudaSuccess) te = cudaEventRecord(e1, stream);
            if (te == cudaSuccess) te = cudaEventSynchronize(e1);
            if (te == cudaSuccess) te = cudaEventElapsedTime(ms, e0, e1);
            if (te != cudaSuccess) (void)cudaGetLastError();
            return te == cudaSuccess;
          };
          float ms_patch = 0.f, ms_tsum = 0.f;
          use = time_it([&] { return launch_conv_patch(out->pp, num_sms_, stream); }, &ms_patch) &&
                time_it([&] { return launch_conv_tsum(ts, num_sms_, stream); }, &ms_tsum) &&
                ms_tsum < ms_patch * 0.97f;
          cudaEventDestroy(e1);
        }
        cudaEventDestroy(e0);
      }
      if (use) {
        out->patch = false;
        out->tsum = true;
        out->ts = ts;
      }
    }
    return true;
  }
  if (!tune) {
    // no timing (a batch size nobody prepared): the tap kernel with the traffic model's cluster
    // shape runs every geometry; host work only
    out->patch = false;
    return build_conv_tc_params(op.g, batch, x, w, y, op.d_shift, res, op.relu, num_sms_, &out->tc,
                                &perr, 0, 0, 0, 0, 0, xe)
               ? true
               : fail(perr);
  }
  cudaEvent_t e0, e1;
  if (!cuda_ok(cudaEventCreate(&e0), "cudaEventCreate", err) ||
      !cuda_ok(cudaEventCreate(&e1), "cudaEventCreate", err))
    return false;
  auto time_it = [&](auto&& launch, float* ms) -> bool {
    cudaError_t te = launch();
    if (te == cudaSuccess) te = cudaEventRecord(e0, stream);
    for (int r = 0; r < 8 && te == cudaSuccess; ++r) te = launch();
    if (te == cudaSuccess) te = cudaEventRecord(e1, stream);
    if (te == cudaSuccess) te = cudaEventSynchronize(e1);
    if (te == cudaSuccess) te = cudaEventElapsedTime(ms, e0, e1);
    if (te != cudaSuccess) perr = std::string("plan timing: ") + cudaGetErrorString(te);
    return te == cudaSuccess;
  };
  bool ok = false, failed = false;
  float best_ms = -1.f;
  if (patch_ok) {
    ConvPatchParams trial;
    float ms = 0.f;
    if (!build_conv_patch_params(op.g, batch, x, w, y, op.d_shift, res, op.relu, &trial, &perr) ||
        !time_it([&] { return launch_conv_patch(trial, num_sms_, stream); }, &ms)) {
      failed = true;
    } else {
      best_ms = ms;
      out->patch = true;
      out->pp = trial;
      ok = true;
    }
  }
  static const int cand[6][2] = {{1, 1}, {2, 1}, {4, 1}, {1, 2}, {2, 2}, {1, 4}};
  for (int ci = 0; ci < 6 && !failed; ++ci) {
    ConvTcParams trial;
    if (!build_conv_tc_params(op.g, batch, x, w, y, op.d_shift, res, op.relu, num_sms_, &trial,
                              &perr, cand[ci][0], cand[ci][1], 0, 0, 0, xe)) {
      failed = true;
      break;
    }
    if (trial.cm != cand[ci][0] || trial.cn != cand[ci][1]) continue;  // shape not applicable
    float ms = 0.f;
    if (!time_it([&] { return launch_conv_tc(trial, num_sms_, stream); }, &ms)) {
      failed = true;
      break;
    }
    // leave the earlier (simpler) candidate only for a clear win
    if (best_ms < 0 || ms < best_ms * 0.97f) {
      best_ms = ms;
      out->patch = false;
      out->tc = trial;
      ok = true;
    }
  }
  // One-wave channel tiling for the tap kernel: when there are fewer position tiles than SMs, a
  // channel tile chosen so that position tiles x channel tiles just fits the SM count avoids a
  // mostly empty second wave (conv4 temporal: 49 x 3 tiles of 96/96/64 channels = 147 CTAs).
  if (!failed && ok) {
    ConvTcParams probe;
    std::string serr;
    if (build_conv_tc_params(op.g, batch, x, w, y, op.d_shift, res, op.relu, num_sms_, &probe, &serr,
                             1, 1, 0, 0, 0, xe) &&
        probe.m_tiles < num_sms_) {
      const int nt = num_sms_ / probe.m_tiles;
      int bn = (op.g.cout_pad() + nt - 1) / nt;
      bn = (bn + 15) / 16 * 16;
      if (bn > kMaxBlockN) bn = kMaxBlockN;
      ConvTcParams trial;
      if (bn >= 32 && bn != probe.block_n &&
          build_conv_tc_params(op.g, batch, x, w, y, op.d_shift, res, op.relu, num_sms_, &trial,
                               &serr, 1, 1, 0, bn, 0, xe)) {
        float ms = 0.f;
        if (!time_it([&] { return launch_conv_tc(trial, num_sms_, stream); }, &ms)) {
          failed = true;
        } else if (ms < best_ms * 0.97f) {
          best_ms = ms;
          out->patch = false;
          out->tc = trial;
        }
      }
    }
  }
  // split-K (conv_splitk.cu) where the tap kernel's grid leaves SMs idle or re-streams the
  // operands of narrow channel tiles: clusters of 2-4 CTAs per output tile, wide or medium tiles.
  // RNB_B200_SPLITK=0 keeps it out of the candidate set.
  static const bool allow_splitk = [] {
    const char* e = getenv("RNB_B200_SPLITK");
    return !(e && e[0] == '0');
  }();
  if (allow_splitk && !failed && ok && !out->patch) {
    // (the 6- and 8-way splits and the 64-wide tiles only ever win at one to three clips, where a
    // conv4 / conv5 layer has a single position tile: per-video latency of the 1-clip videos)
    static const int sk_cand[10][2] = {{2, 256}, {3, 256}, {2, 128}, {3, 128}, {4, 128},
                                       {6, 128}, {8, 128}, {2, 64},  {4, 64},  {8, 64}};
    const int tap_items = out->tc.m_tiles * out->tc.n_tiles;
    for (int ci = 0; ci < 10 && !failed; ++ci) {
      ConvTcParams trial;
      std::string serr;
      if (!build_conv_tc_params(op.g, batch, x, w, y, op.d_shift, res, op.relu, num_sms_, &trial,
                                &serr, 0, 0, sk_cand[ci][0], sk_cand[ci][1], 0, xe))
        continue;  // e.g. fewer K blocks than CTAs
      // only layers short of work items: more than ~3 waves of clusters is the tap kernel's domain
      if (trial.m_tiles * trial.n_tiles * trial.ks > 3 * num_sms_ && tap_items > 2 * num_sms_)
        continue;
      float ms = 0.f;
      if (!time_it([&] { return launch_conv_splitk(trial, num_sms_, stream); }, &ms)) {
        failed = true;
        break;
      }
      if (ms < best_ms * 0.97f) {
        best_ms = ms;
        out->patch = false;
        out->tc = trial;
      }
    }
  }
  // CTA pairs (conv_pair.cu): half the weight bytes per SM, the larger operand of the conv3-conv5
  // layers; channel tiles of at most 256 / 192 / 128 / 64.  RNB_B200_PAIR=0 keeps them out.
  static const bool allow_pair = [] {
    const char* e = getenv("RNB_B200_PAIR");
    return !(e && e[0] == '0');
  }();
  if (allow_pair && !failed && ok) {
    int seen[4] = {0, 0, 0, 0};
    static const int caps[4] = {256, 192, 128, 64};
    for (int ci = 0; ci < 4 && !failed; ++ci) {
      ConvTcParams trial;
      std::string serr;
      if (!build_conv_tc_params(op.g, batch, x, w, y, op.d_shift, res, op.relu, num_sms_, &trial,
                                &serr, 0, 0, 0, caps[ci], 1, xe))
        continue;  // e.g. a single position tile
      bool dup = false;
      for (int k = 0; k < ci; ++k) dup = dup || seen[k] == trial.block_n;
      seen[ci] = trial.block_n;
      if (dup) continue;
      float ms = 0.f;
      if (!time_it([&] { return launch_conv_pair(trial, num_sms_, stream); }, &ms)) {
        failed = true;
        break;
      }
      if (ms < best_ms * 0.97f) {
        best_ms = ms;
        out->patch = false;
        out->tc = trial;
      }
    }
  }
  // temporal-patch mode of the CTA-pair kernel: wide 3x1x1 stride-1 layers (conv3 temporal).
  // Opt-in (RNB_B200_TPATCH=1): 15.3 -> 12.9 us on conv3's temporal layers in isolation but no
  // measurable change of the step, and its K order (block-outer, tap-inner) differs from the other
  // kernels', which costs the bit-identical rows across batch sizes.
  static const bool allow_tpatch = [] {
    const char* e = getenv("RNB_B200_TPATCH");
    return e && e[0] == '1';
  }();
  if (allow_tpatch && allow_pair && !failed && ok && conv_tpatch_applicable(op.g)) {
    int seen_n = 0;
    for (int cap : {256, 128}) {
      ConvTcParams trial;
      std::string serr;
      if (!build_conv_tpatch_params(op.g, batch, x, w, y, op.d_shift, res, op.relu, cap, &trial,
                                    &serr))
        continue;
      if (trial.block_n == seen_n) continue;
      seen_n = trial.block_n;
      float ms = 0.f;
      if (!time_it([&] { return launch_conv_pair(trial, num_sms_, stream); }, &ms)) {
        failed = true;
        break;
      }
      if (ms < best_ms * 0.97f) {
        best_ms = ms;
        out->patch = false;
        out->tc = trial;
      }
    }
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (failed || !ok) return fail(perr.empty() ? "no launch candidate" : perr);
  return true;
}

// The two stem entries of a plan in fused mode: entry 0 runs both convolutions, entry 1 is merged.
bool R2P1DModel::plan_stem(int batch, std::vector<Launch>* plan, std::string* err) {
  Launch a, b;
  a.op = 0;
  a.batch = batch;
  a.stem_fused = true;
  if (!build_stem_fused_params(batch, (const __nv_bfloat16*)slots_[SLOT_U],
                               (const __nv_bfloat16*)d_stem_w_, (const __nv_bfloat16*)ops_[1].d_w,
                               (__nv_bfloat16*)slots_[ops_[1].out_slot], &a.sf, err))
    return false;
  b.op = 1;
  b.batch = batch;
  b.merged = true;
  plan->push_back(a);
  plan->push_back(b);
  return true;
}

// Entry i of a plan as one launch (nothing for an entry merged into its predecessor).
cudaError_t R2P1DModel::launch_one(const std::vector<Launch>& plan, size_t i,
                                   cudaStream_t stream) const {
  const Launch& l = plan[i];
  if (l.merged) return cudaSuccess;
  if (l.stem_fused) return launch_conv_stem_fused(l.sf, num_sms_, stream);
  if (l.dual_n_sp > 0 && i + 1 < plan.size())
    return launch_conv_patch_dual(l.pp, plan[i + 1].pp, l.dual_n_sp, l.dual_chunk, d_dual_flags_,
                                  max_batch_, num_sms_, stream);
  if (l.tsum) return launch_conv_tsum(l.ts, num_sms_, stream);
  return l.patch ? launch_conv_patch(l.pp, num_sms_, stream) : launch_tc(l.tc, stream);
}

// Entries i, i+1 = the two convolutions of one SpatioTemporalConv, both on the resident-weight
// patch kernel over the whole batch: the pair the co-scheduled kernel can run in one launch.
bool R2P1DModel::dual_candidate(const std::vector<Launch>& plan, int i) const {
  if (i + 1 >= (int)plan.size()) return false;
  const Launch& a = plan[i];
  const Launch& b = plan[i + 1];
  if (!a.patch || !b.patch || a.op + 1 != b.op) return false;
  if (ops_[a.op].out_slot != SLOT_M || ops_[b.op].in_slot != SLOT_M) return false;
  if (a.first != 0 || b.first != 0 || a.batch != b.batch) return false;
  return conv_patch_dual_applicable(a.pp, b.pp);
}

// Times the pair as two launches and as one co-scheduled launch at a few CTA splits; keeps the
// co-scheduled form when it is clearly faster (it reads the intermediate tensor from L2, but the
// two convolutions then share the SMs' bandwidth and power: it pays where BOTH are bandwidth-bound,
// as in the stem, not where the spatial one is tensor-bound, as in conv2;
// profiles/r02_coscheduled_conv2.md).  RNB_B200_DUAL=0 keeps it out.
bool R2P1DModel::tune_dual(std::vector<Launch>* plan, int i, cudaStream_t stream) {
  static const bool allow = [] {
    const char* e = getenv("RNB_B200_DUAL");
    return !(e && e[0] == '0');
  }();
  if (!allow || !dual_candidate(*plan, i)) return false;
  Launch& a = (*plan)[i];
  Launch& b = (*plan)[i + 1];
  cudaEvent_t e0, e1;
  if (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess) return false;
  auto time_it = [&](auto&& launch, float* ms) -> bool {
    cudaError_t te = launch();
    if (te == cudaSuccess) te = cudaEventRecord(e0, stream);
    for (int r = 0; r < 8 && te == cudaSuccess; ++r) te = launch();
    if (te == cudaSuccess) te = cudaEventRecord(e1, stream);
    if (te == cudaSuccess) te = cudaEventSynchronize(e1);
    if (te == cudaSuccess) te = cudaEventElapsedTime(ms, e0, e1);
    if (te != cudaSuccess) (void)cudaGetLastError();
    return te == cudaSuccess;
  };
  float base = 0.f;
  bool ok = time_it([&] {
    cudaError_t e = launch_conv_patch(a.pp, num_sms_, stream);
    return e == cudaSuccess ? launch_conv_patch(b.pp, num_sms_, stream) : e;
  }, &base);
  float best = base;
  int best_n = 0;
  static const int kPct[] = {50, 58, 66, 74, 82};
  for (int pct : kPct) {
    if (!ok) break;
    const int n_sp = num_sms_ * pct / 100;
    if (n_sp < 1 || n_sp >= num_sms_) continue;
    float ms = 0.f;
    ok = time_it([&] {
      return launch_conv_patch_dual(a.pp, b.pp, n_sp, 4, d_dual_flags_, max_batch_, num_sms_, stream);
    }, &ms);
    if (ok && ms < best * 0.97f) {
      best = ms;
      best_n = n_sp;
    }
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (!ok || best_n == 0) return false;
  a.dual_n_sp = best_n;
  a.dual_chunk = 4;
  b.merged = true;
  return true;
}

cudaError_t R2P1DModel::launch_tc(const ConvTcParams& tc, cudaStream_t stream) const {
  if (tc.pair) return launch_conv_pair(tc, num_sms_, stream);
  return tc.ks > 1 ? launch_conv_splitk(tc, num_sms_, stream) : launch_conv_tc(tc, num_sms_, stream);
}

// The flat launch list of one forward pass at `batch` clips.  Optionally (RNB_B200_CHUNK_MB > 0) a
// SpatioTemporalConv whose intermediate tensor (spatial output = temporal input, slot M) does not
// survive in L2 is run in batch chunks -- spatial(chunk), temporal(chunk), ... -- so that the
// intermediate is produced and consumed while cache resident, reusing the first chunk-sized part
// of slot M.  ncu shows the unchunked conv2 temporal kernel re-reading all 115 MB from DRAM
// (profiles/r01_ncu_conv_kernels_v2.md), but measured on B200 at B=16 the chunked schedule is
// SLOWER (15.7k vs 16.8k clips/s at 32 MB chunks, 12.9k at 16 MB): every extra launch reloads the
// resident weights and runs a short, badly quantised grid.  Hence off by default.
bool R2P1DModel::build_plan(int batch, cudaStream_t stream, std::vector<Launch>* plan,
                            std::string* err) {
  static const double chunk_mb = [] {
    const char* e = getenv("RNB_B200_CHUNK_MB");
    return e ? atof(e) : 0.0;
  }();
  const int n = (int)ops_.size();
  if (stem_fused_ && !plan_stem(batch, plan, err)) return false;
  for (int i = stem_fused_ ? 2 : 0; i < n; ++i) {
    const ConvOp& op = ops_[i];
    const bool pair = i + 1 < n && op.out_slot == SLOT_M && ops_[i + 1].in_slot == SLOT_M;
    const double mid_mb_per_clip = (double)op.g.out_elems(1) * 2 / 1e6;
    int bc = chunk_mb > 0 ? (int)(chunk_mb / mid_mb_per_clip) : batch;
    if (bc < 1) bc = 1;
    if (!pair || mid_mb_per_clip * batch <= 1.5 * chunk_mb || bc >= batch) {
      Launch l;
      if (!plan_conv(i, 0, batch, false, stream, nullptr, &l, err)) return false;
      plan->push_back(l);
      continue;
    }
    // chunked pair: tune once per distinct chunk size, copy the choice to the other chunks
    std::map<int, std::pair<Launch, Launch>> tuned;
    for (int first = 0; first < batch; first += bc) {
      const int b = batch - first < bc ? batch - first : bc;
      auto it = tuned.find(b);
      Launch sp, tp;
      if (!plan_conv(i, first, b, true, stream, it == tuned.end() ? nullptr : &it->second.first,
                     &sp, err) ||
          !plan_conv(i + 1, first, b, true, stream,
                     it == tuned.end() ? nullptr : &it->second.second, &tp, err))
        return false;
      if (it == tuned.end()) tuned.emplace(b, std::make_pair(sp, tp));
      plan->push_back(sp);
      plan->push_back(tp);
    }
    ++i;  // the temporal half is planned
  }
  if (plan->size() == ops_.size())
    for (int i = 0; i + 1 < n; ++i)
      if (tune_dual(plan, i, stream)) ++i;
  return true;
}

// Plan cache (RNB_B200_PLAN_CACHE=<file>): the per-layer kernel choices of a batch size are appended
// to the file after they have been timed, and reused -- without timing -- by any later process
// whose model and batch size match.  Makes runs reproducible (the timing picks can differ between
// near-equal candidates) and lets a profiler run use the plan of an unprofiled one.
namespace {
uint64_t fnv1a(uint64_t h, const void* data, size_t n) {
  const unsigned char* p = static_cast<const unsigned char*>(data);
  for (size_t i = 0; i < n; ++i) h = (h ^ p[i]) * 1099511628211ull;
  return h;
}
}  // namespace

uint64_t R2P1DModel::plan_key(int batch) const {
  uint64_t h = 1469598103934665603ull;
  h = fnv1a(h, &batch, sizeof batch);
  h = fnv1a(h, &num_sms_, sizeof num_sms_);
  for (const ConvOp& op : ops_) {
    h = fnv1a(h, op.name.data(), op.name.size());
    const int g[8] = {op.g.Ti, op.g.Hi, op.g.Wi, op.g.cin, op.g.cout, op.g.extra_cin, op.g.ncat,
                      op.res_slot >= 0};
    h = fnv1a(h, g, sizeof g);
  }
  return h;
}

bool R2P1DModel::load_cached_plan(const char* path, int batch, cudaStream_t stream,
                                  std::vector<Launch>* plan) {
  FILE* f = fopen(path, "r");
  if (!f) return false;
  const unsigned long long want = plan_key(batch);
  char line[256];
  bool found = false;
  while (fgets(line, sizeof line, f)) {
    unsigned long long key = 0;
    int n = 0;
    if (sscanf(line, "plan %llx %d", &key, &n) != 2 || key != want || n != (int)ops_.size()) continue;
    found = true;
    for (int i = 0; i < n && found; ++i) {
      int op = -1, patch = 0, cm = 1, cn = 1, ks = 0, bn = 0, pair = 0, tpatch = 0, dual = 0, tsum = 0;
      if (!fgets(line, sizeof line, f) ||
          sscanf(line, "%d %d %d %d %d %d %d %d %d %d", &op, &patch, &cm, &cn, &ks, &bn, &pair,
                 &tpatch, &dual, &tsum) < 7 ||
          op != i) {
        found = false;
        break;
      }
      if (stem_fused_ && i < 2) {
        if (i == 0) {
          std::string e;
          if (!plan_stem(batch, plan, &e)) found = false;
        }
        continue;
      }
      Launch like, l;
      like.patch = patch != 0;
      like.tc.cm = cm;
      like.tc.cn = cn;
      like.tc.ks = ks;
      like.tc.block_n = bn;
      like.tc.pair = pair;
      like.pp.pair = pair;
      like.tc.tpatch = tpatch;
      like.tsum = tsum != 0;
      std::string e;
      if (!plan_conv(i, 0, batch, false, stream, &like, &l, &e)) found = false;
      l.dual_n_sp = dual;
      l.dual_chunk = dual > 0 ? 4 : 0;
      plan->push_back(l);
    }
    for (int i = 0; found && i + 1 < n; ++i)
      if ((*plan)[i].dual_n_sp > 0) {
        if (dual_candidate(*plan, i))
          (*plan)[i + 1].merged = true;
        else
          (*plan)[i].dual_n_sp = 0;
      }
    break;
  }
  fclose(f);
  if (!found) plan->clear();
  return found;
}

void R2P1DModel::save_cached_plan(const char* path, int batch, const std::vector<Launch>& plan) {
  if (plan.size() != ops_.size()) return;  // chunked schedules are not cached
  FILE* f = fopen(path, "a");
  if (!f) return;
  fprintf(f, "plan %llx %d\n", (unsigned long long)plan_key(batch), (int)plan.size());
  for (const Launch& l : plan)
    fprintf(f, "%d %d %d %d %d %d %d %d %d %d\n", l.op, l.patch ? 1 : 0, l.tc.cm, l.tc.cn, l.tc.ks,
            l.tc.block_n, l.patch ? l.pp.pair : l.tc.pair, l.tc.tpatch, l.dual_n_sp, l.tsum ? 1 : 0);
  fclose(f);
}

// The plan of a batch size nobody prepared, built WITHOUT timing and without touching the GPU: every
// layer copies the kernel choice of the nearest prepared batch size (cluster shapes and split-K
// factors that do not apply at this size fall back to the plain tap kernel); with nothing prepared
// yet, the patch kernel where it keeps its weights resident and the tap kernel elsewhere.
bool R2P1DModel::derive_plan(int batch, std::vector<Launch>* plan, std::string* err) {
  const std::vector<Launch>* src = nullptr;
  double best = 0;
  for (int b : tuned_) {
    const auto it = plans_.find(b);
    if (it == plans_.end() || it->second.size() != ops_.size()) continue;  // chunked schedules
    const double d = fabs(log((double)b / (double)batch));
    if (!src || d < best) {
      src = &it->second;
      best = d;
    }
  }
  if (stem_fused_ && !plan_stem(batch, plan, err)) return false;
  for (int i = stem_fused_ ? 2 : 0; i < (int)ops_.size(); ++i) {
    Launch l;
    std::string e;
    if (!(src && plan_conv(i, 0, batch, false, nullptr, &(*src)[i], &l, &e)) &&
        !plan_conv(i, 0, batch, false, nullptr, nullptr, &l, err, false))
      return false;
    plan->push_back(l);
  }
  if (src)
    for (int i = 0; i + 1 < (int)ops_.size(); ++i)
      if ((*src)[i].dual_n_sp > 0 && dual_candidate(*plan, i)) {
        (*plan)[i].dual_n_sp = (*src)[i].dual_n_sp;
        (*plan)[i].dual_chunk = (*src)[i].dual_chunk;
        (*plan)[i + 1].merged = true;
      }
  return true;
}

const std::vector<R2P1DModel::Launch>* R2P1DModel::ensure_plan(int batch, cudaStream_t stream,
                                                               std::string* err, bool tune) {
  auto it = plans_.find(batch);
  if (it != plans_.end() && (!tune || tuned_.count(batch))) return &it->second;
  std::vector<Launch> plan;
  if (tune) {
    const char* cache = getenv("RNB_B200_PLAN_CACHE");
    if (!(cache && cache[0] && load_cached_plan(cache, batch, stream, &plan))) {
      if (!build_plan(batch, stream, &plan, err)) return nullptr;
      if (cache && cache[0]) save_cached_plan(cache, batch, plan);
    }
    tuned_.insert(batch);
  } else if (!derive_plan(batch, &plan, err)) {
    return nullptr;
  }
  set_clip_order(&plan);
  plans_[batch] = std::move(plan);
  return &plans_[batch];
}

// Alternating clip order (ConvPatchParams::reverse): the 1 x k x k convolutions take the clips last
// to first, the k x 1 x 1 convolutions (and the fused stem) first to last, so every patch launch
// starts on the clips its predecessor wrote last, which are the ones still in L2.
// RNB_B200_CLIP_ORDER=forward keeps every launch first-to-last.
void R2P1DModel::set_clip_order(std::vector<Launch>* plan) const {
  static const bool snake = [] {
    const char* e = getenv("RNB_B200_CLIP_ORDER");
    return !(e && e[0] == 'f');
  }();
  for (Launch& l : *plan)
    if (l.patch && !l.stem_fused)
      l.pp.reverse = snake && l.dual_n_sp == 0 && !l.merged && ops_[l.op].g.kh > 1 ? 1 : 0;
}

bool R2P1DModel::plan_entry(int batch, int op, int out[7]) const {
  const auto it = plans_.find(batch);
  if (it == plans_.end()) return false;
  for (const Launch& l : it->second)
    if (l.op == op) {
      out[0] = l.patch ? 2 : (l.tc.pair ? (l.tc.tpatch ? 5 : 4) : (l.tc.ks > 1 ? 3 : 1));
      if (l.tsum) out[0] = 8;
      if (l.dual_n_sp > 0 || l.merged) out[0] = 6;
      if (l.stem_fused || (stem_fused_ && op < 2)) out[0] = 7;
      const bool one = l.patch || l.tsum;  // single-CTA kernels without cluster / split-K shape
      out[1] = l.dual_n_sp > 0 ? l.dual_n_sp : (one ? 1 : l.tc.cm);
      out[2] = one ? 1 : l.tc.cn;
      out[3] = one ? 0 : l.tc.ks;
      out[4] = l.tsum ? kTsumN : (l.patch ? l.pp.block_n : l.tc.block_n);
      out[5] = l.tsum ? 0 : (l.patch ? l.pp.pair : l.tc.pair);
      out[6] = tuned_.count(batch) ? 1 : 0;
      return true;
    }
  return false;
}

int R2P1DModel::prepare(int batch, cudaStream_t stream, std::string* err) {
  if (!finalized_) {
    if (err) *err = "prepare before finalize";
    return RNB_ESTATE;
  }
  if (batch < 1 || batch > max_batch_) {
    if (err)
      *err = "batch " + std::to_string(batch) + " outside [1, max_batch = " +
             std::to_string(max_batch_) + "]";
    return RNB_EINVAL;
  }
  if (dtype_ != RNB_DTYPE_BF16) return RNB_OK;  // the fp32 parity mode has one kernel
  return ensure_plan(batch, stream, err, true) ? RNB_OK : RNB_ECUDA;
}

int R2P1DModel::force_kernel(int batch, int op_index, int kind, int cm, int cn, int ks, int block_n,
                             int pair, std::string* err) {
  if (!finalized_ || dtype_ != RNB_DTYPE_BF16) {
    if (err) *err = "force_kernel needs a finalized bf16 handle";
    return RNB_ESTATE;
  }
  if (batch < 1 || batch > max_batch_ || op_index < 0 || op_index >= (int)ops_.size()) {
    if (err) *err = "force_kernel: batch or op index out of range";
    return RNB_EINVAL;
  }
  if (stem_fused_ && op_index < 2) {
    if (err) *err = "the stem runs as one fused kernel (RNB_B200_STEM=unfolded to pin its halves)";
    return RNB_EINVAL;
  }
  if (!ensure_plan(batch, nullptr, err, false)) return RNB_ECUDA;
  std::vector<Launch>& plan = plans_[batch];
  if (plan.size() != ops_.size()) {
    if (err) *err = "force_kernel: chunked schedules cannot be edited";
    return RNB_ESTATE;
  }
  if (kind == 6) {
    // co-scheduled launch of this spatial convolution and the temporal one that follows; cm = the
    // CTAs that start on the spatial convolution (0: three quarters of the SMs)
    std::string e;
    Launch a, b;
    Launch lp;
    lp.patch = true;
    if (op_index + 1 >= (int)ops_.size() || !plan_conv(op_index, 0, batch, false, nullptr, &lp, &a, &e) ||
        !plan_conv(op_index + 1, 0, batch, false, nullptr, &lp, &b, &e)) {
      if (err) *err = ops_[op_index].name + ": the co-scheduled kernel does not apply here";
      return RNB_EINVAL;
    }
    std::vector<Launch> probe = plan;
    probe[op_index] = a;
    probe[op_index + 1] = b;
    if (!dual_candidate(probe, op_index)) {
      if (err) *err = ops_[op_index].name + ": the co-scheduled kernel does not apply here";
      return RNB_EINVAL;
    }
    a.dual_n_sp = cm > 0 && cm < num_sms_ ? cm : num_sms_ * 3 / 4;
    a.dual_chunk = cn > 0 ? cn : 4;
    b.merged = true;
    plan[op_index] = a;
    plan[op_index + 1] = b;
    set_clip_order(&plan);
    return RNB_OK;
  }
  // (pinning one half of a co-scheduled pair on its own dissolves the pair)
  if (plan[op_index].dual_n_sp > 0 && op_index + 1 < (int)plan.size()) plan[op_index + 1].merged = false;
  if (plan[op_index].merged && op_index > 0) {
    plan[op_index - 1].dual_n_sp = 0;
    plan[op_index].merged = false;
  }
  Launch like, l;
  switch (kind) {
    case 1: like.tc.cm = cm > 0 ? cm : 1; like.tc.cn = cn > 0 ? cn : 1; like.tc.block_n = block_n; break;
    case 2: like.patch = true; like.pp.pair = pair ? 1 : 0; break;
    case 3: like.tc.ks = ks; like.tc.block_n = block_n; break;
    case 4: like.tc.pair = 1; like.tc.block_n = block_n; break;
    case 5: like.tc.pair = 1; like.tc.tpatch = 1; like.tc.block_n = block_n; break;
    case 8: like.tsum = true; break;
    default:
      if (err) *err = "force_kernel: unknown kernel kind " + std::to_string(kind);
      return RNB_EINVAL;
  }
  if (!plan_conv(op_index, 0, batch, false, nullptr, &like, &l, err)) return RNB_EINVAL;
  // the builders fall back silently where a shape does not apply: report that as an error here
  const bool ok = kind == 1   ? (!l.patch && l.tc.cm == like.tc.cm && l.tc.cn == like.tc.cn &&
                                 l.tc.ks <= 1 && !l.tc.pair)
                  : kind == 2 ? (l.patch && l.pp.pair == like.pp.pair)
                  : kind == 3 ? (!l.patch && l.tc.ks == ks && ks > 1)
                  : kind == 4 ? (!l.patch && l.tc.pair && !l.tc.tpatch)
                  : kind == 8 ? l.tsum
                              : (!l.patch && l.tc.pair && l.tc.tpatch);
  if (!ok) {
    if (err) *err = ops_[op_index].name + ": the requested kernel shape does not apply to this layer";
    return RNB_EINVAL;
  }
  plan[op_index] = l;
  set_clip_order(&plan);
  return RNB_OK;
}

int R2P1DModel::forward(const void* in, int in_fmt, int batch, void* out, int out_fmt,
                        cudaStream_t stream, std::string* err, cudaEvent_t* events,
                        float* video_sum, int* video_cls, bool video) {
  int ev_idx = 0;
  auto mark = [&]() {
    if (events) cudaEventRecord(events[ev_idx++], stream);
  };
  if (!finalized_) {
    if (err) *err = "forward before finalize";
    return RNB_ESTATE;
  }
  if (batch < 0 || batch > max_batch_) {
    if (err)
      *err = "batch " + std::to_string(batch) + " outside [0, max_batch = " +
             std::to_string(max_batch_) + "]";
    return RNB_EINVAL;
  }
  if (batch == 0) return RNB_OK;
  if (in == nullptr || (out == nullptr && !video)) {
    if (err) *err = "null input or output pointer";
    return RNB_EINVAL;
  }
  if (video && end_idx_ != 5) {
    if (err) *err = "per-video classification needs a model that ends at layer 5";
    return RNB_EINVAL;
  }
  const bool tc = dtype_ == RNB_DTYPE_BF16;
  const int want_out = end_idx_ == 5 ? RNB_FMT_F32_LOGITS : out_fmt;
  if (out_fmt != want_out ||
      (end_idx_ != 5 && out_fmt != RNB_FMT_F32_NCDHW && !(tc && out_fmt == RNB_FMT_BF16_NDHWC))) {
    if (err) *err = "unsupported output format " + std::to_string(out_fmt) + " for this handle";
    return RNB_EINVAL;
  }
  // ---- launch plan for this batch size: what prepare() timed, or -- for a size nobody prepared --
  // a plan derived on the host from the nearest prepared size (no timing, no synchronisation).
  // (Every TMA tensor map of a plan covers only this handle's own arena and weights, allocated on
  // this device; caller pointers are read by the staging kernels / copies below, never by TMA.)
  const std::vector<Launch>* plan = nullptr;
  if (tc) {
    plan = ensure_plan(batch, stream, err, false);
    last_batch_ = batch;
    if (!plan) return RNB_ECUDA;
    // L2 persistence carve-out on or off for this batch size (conv_tc.cuh)
    size_t largest = 0;
    for (const ConvOp& op : ops_) {
      const size_t b = (size_t)op.g.out_elems(batch) * 2;
      if (b > largest) largest = b;
    }
    l2_persist_configure(largest);
  }
  cudaError_t e = cudaSuccess;
  mark();

  // ---- stage the input into the arena
  const int S_in = in_dims_.T * in_dims_.H * in_dims_.W;
  if (start_idx_ == 1) {
    if (in_fmt == RNB_FMT_U8_NTCHW) {
      e = tc ? (stem_fused_ ? launch_pre_u8_to_s2d((const uint8_t*)in, batch, nullptr, nullptr,
                                                   (__nv_bfloat16*)slots_[SLOT_U], stream)
                            : launch_pre_u8_to_stem((const uint8_t*)in, batch, nullptr, nullptr,
                                                    (__nv_bfloat16*)slots_[SLOT_U], stream))
             : launch_pre_u8_to_f32_ndhwc((const uint8_t*)in, batch, nullptr, nullptr,
                                          (float*)slots_[SLOT_U], stream);
    } else if (in_fmt == RNB_FMT_F32_NCDHW) {
      e = tc ? (stem_fused_ ? launch_f32_ncdhw_to_s2d((const float*)in, batch,
                                                      (__nv_bfloat16*)slots_[SLOT_U], stream)
                            : launch_f32_ncdhw_to_stem((const float*)in, batch,
                                                       (__nv_bfloat16*)slots_[SLOT_U], stream))
             : launch_ncdhw_f32_to_ndhwc_f32((const float*)in, batch, kClipC, S_in,
                                             (float*)slots_[SLOT_U], stream);
    } else {
      if (err) *err = "a model starting at layer 1 takes U8_NTCHW or F32_NCDHW input";
      return RNB_EINVAL;
    }
  } else {
    if (in_fmt == RNB_FMT_F32_NCDHW) {
      e = tc ? launch_ncdhw_f32_to_ndhwc_bf16((const float*)in, batch, in_dims_.C, S_in,
                                              pad16(in_dims_.C), (__nv_bfloat16*)slots_[SLOT_X],
                                              stream)
             : launch_ncdhw_f32_to_ndhwc_f32((const float*)in, batch, in_dims_.C, S_in,
                                             (float*)slots_[SLOT_X], stream);
    } else if (in_fmt == RNB_FMT_BF16_NDHWC && tc) {
      e = cudaMemcpyAsync(slots_[SLOT_X], in, (size_t)batch * S_in * pad16(in_dims_.C) * 2,
                          cudaMemcpyDeviceToDevice, stream);
    } else {
      if (err) *err = "unsupported input format " + std::to_string(in_fmt) + " for this handle";
      return RNB_EINVAL;
    }
  }
  if (!cuda_ok(e, "input staging", err)) return RNB_ECUDA;
  mark();

  // ---- convolutions
  if (tc) {
    for (size_t i = 0; i < plan->size(); ++i) {
      const Launch& l = (*plan)[i];
      if (l.merged) continue;  // runs inside the previous, co-scheduled launch
      if (!cuda_ok(launch_one(*plan, i, stream), ops_[l.op].name.c_str(), err)) return RNB_ECUDA;
      mark();
    }
  } else {
    for (const ConvOp& op : ops_) {
      ConvRefParams rp;
      rp.x = (const float*)slots_[op.in_slot];
      rp.w = (const float*)op.d_w;
      rp.scale = op.d_scale;
      rp.shift = op.d_shift;
      rp.residual = op.res_slot >= 0 ? (const float*)slots_[op.res_slot] : nullptr;
      rp.y = (float*)slots_[op.out_slot];
      rp.B = batch; rp.Ti = op.g.Ti; rp.Hi = op.g.Hi; rp.Wi = op.g.Wi; rp.Cin = op.g.cin;
      rp.To = op.g.To; rp.Ho = op.g.Ho; rp.Wo = op.g.Wo; rp.Cout = op.g.cout;
      rp.kt = op.g.kt; rp.kh = op.g.kh; rp.kw = op.g.kw;
      rp.st = op.g.st; rp.sh = op.g.sh; rp.sw = op.g.sw;
      rp.pt = op.g.pt; rp.ph = op.g.ph; rp.pw = op.g.pw;
      rp.relu = op.relu ? 1 : 0;
      if (!cuda_ok(launch_conv_ref(rp, stream), op.name.c_str(), err)) return RNB_ECUDA;
      mark();
    }
  }

  // ---- output
  const int S_out = out_dims_.T * out_dims_.H * out_dims_.W;
  if (end_idx_ == 5 && video && tc) {
    e = launch_pool_fc_classify_bf16((const __nv_bfloat16*)slots_[final_slot_], batch, S_out, 512,
                                     d_fc_w_, d_fc_b_, num_classes_, (float*)out, video_sum,
                                     video_cls, d_classify_scratch_, stream);
  } else if (end_idx_ == 5 && video) {
    // fp32 parity mode: separate classifier and reduction kernels, logits staged in the scratch
    float* lg = out ? (float*)out : reinterpret_cast<float*>((char*)d_classify_scratch_ + 256);
    e = launch_pool_fc_f32((const float*)slots_[final_slot_], batch, S_out, 512, d_fc_w_, d_fc_b_,
                           num_classes_, lg, stream);
    if (e == cudaSuccess)
      e = launch_logits_sum_argmax(lg, batch, num_classes_, video_sum, video_cls, stream);
  } else if (end_idx_ == 5) {
    e = tc ? launch_pool_fc_bf16((const __nv_bfloat16*)slots_[final_slot_], batch, S_out, 512,
                                 d_fc_w_, d_fc_b_, num_classes_, (float*)out, stream)
           : launch_pool_fc_f32((const float*)slots_[final_slot_], batch, S_out, 512, d_fc_w_,
                                d_fc_b_, num_classes_, (float*)out, stream);
  } else if (out_fmt == RNB_FMT_F32_NCDHW) {
    e = tc ? launch_ndhwc_bf16_to_ncdhw_f32((const __nv_bfloat16*)slots_[final_slot_], batch,
                                            out_dims_.C, S_out, pad16(out_dims_.C), (float*)out,
                                            stream)
           : launch_ndhwc_f32_to_ncdhw_f32((const float*)slots_[final_slot_], batch, out_dims_.C,
                                           S_out, (float*)out, stream);
  } else {
    e = cudaMemcpyAsync(out, slots_[final_slot_], (size_t)batch * S_out * pad16(out_dims_.C) * 2,
                        cudaMemcpyDeviceToDevice, stream);
  }
  if (!cuda_ok(e, "output stage", err)) return RNB_ECUDA;
  mark();
  last_launches_ = 2;
  if (tc) {
    for (const Launch& l : *plan) last_launches_ += l.merged ? 0 : 1;
  } else {
    last_launches_ += (int)ops_.size();
  }
  return RNB_OK;
}

}  // namespace rnb
