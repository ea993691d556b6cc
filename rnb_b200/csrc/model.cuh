// R(2+1)D layer-range model: op list, folded parameters, activation arena, per-batch launch plans.
#pragma once
#include <map>
#include <set>
#include <string>
#include <vector>

#include "conv_tc.cuh"
#include "kernels.cuh"
#include "plan.cuh"

namespace rnb {

struct HostTensor {
  std::vector<int64_t> shape;
  std::vector<float> data;
};

// activation slots of the arena; every op reads one or two slots and writes a third
// U stem input, X/Y block in/out, M intermediate of a factorised conv, R first conv output,
// S shortcut output (fp32 mode), D shortcut intermediate (bf16 mode: the shortcut's temporal
// convolution is fused into the block's last convolution and reads it from there)
enum Slot { SLOT_U = 0, SLOT_X, SLOT_M, SLOT_R, SLOT_S, SLOT_Y, SLOT_D, NUM_SLOTS };

struct ConvOp {
  std::string name;  // state_dict prefix of the Conv3d, for error messages
  ConvGeom g;
  int in_slot = 0, out_slot = 0, res_slot = -1;
  int extra_slot = -1;  // second input of a fused shortcut tap (ConvGeom::extra_*)
  bool relu = false;
  std::vector<float> w;             // torch layout (cout, cin, kt, kh, kw)
  std::vector<float> scale, shift;  // folded BatchNorm + bias, length cout
  std::vector<float> w_extra, scale_extra;  // K-concatenated shortcut convolution (cout, extra_cin)
  std::vector<float> w_ncat, scale_ncat;    // N-concatenated shortcut convolution (ncat, cin)
  // device copies
  void* d_w = nullptr;  // bf16 K-major rows (tensor cores) or fp32 (kt,kh,kw,cin,cout)
  float* d_scale = nullptr;
  float* d_shift = nullptr;
};

class R2P1DModel {
 public:
  R2P1DModel(std::vector<int> layer_sizes, int start_idx, int end_idx, int num_classes, int dtype,
             int max_batch);
  ~R2P1DModel();

  bool load_param(const char* key, const float* data, const int64_t* shape, int ndim,
                  std::string* err);
  int finalize(std::string* err);  // RNB_* code
  // `events`, when given, receives one event before the first kernel and one after every kernel
  int forward(const void* in, int in_fmt, int batch, void* out, int out_fmt, cudaStream_t stream,
              std::string* err, cudaEvent_t* events = nullptr, float* video_sum = nullptr,
              int* video_cls = nullptr, bool video = false);
  int profile(const void* in, int in_fmt, int batch, void* out, int out_fmt, cudaStream_t stream,
              float* ms, int cap, std::string* err);
  // Times the candidate kernels of every layer at `batch` clips and keeps the fastest (the launch
  // plan of that batch size).  Synchronises `stream` and overwrites the activation arena.
  int prepare(int batch, cudaStream_t stream, std::string* err);
  // Pins the kernel of convolution `op` in the plan of `batch` (parity tests, sweeps): kind as in
  // op_info (1 tap, 2 patch, 3 split-K, 4 CTA pair, 5 CTA-pair temporal patch).
  int force_kernel(int batch, int op, int kind, int cm, int cn, int ks, int block_n, int pair,
                   std::string* err);
  bool plan_is_tuned(int batch) const { return tuned_.count(batch) != 0; }
  // The plan entry of convolution `op` at `batch`: {kind, cm, cn, ks, block_n, pair, tuned}.
  bool plan_entry(int batch, int op, int out[7]) const;
  bool op_info(int op, std::string* name, double* flops_per_clip, double* act_bytes_per_clip,
               double* weight_bytes, int* kernel_kind) const;
  void io_shape(int which, int64_t shape4[4]) const;
  double flops(int batch) const;
  // kernels the most recent forward launched (convolutions + staging + classifier/layout)
  int num_launches() const { return last_launches_ > 0 ? last_launches_ : (int)ops_.size() + 2; }
  int num_ops() const { return (int)ops_.size(); }
  bool finalized() const { return finalized_; }

 private:
  struct Dims {
    int C, T, H, W;
  };
  bool build_ops(std::string* err);
  bool add_st_conv(const std::string& prefix, int in_slot, int out_slot, Dims in, int cout, int kt,
                   int kh, int kw, int st, int sh, int sw, int pt, int ph, int pw,
                   const std::string& post_bn, bool relu, int res_slot, bool stem, Dims* out,
                   std::string* err);
  const HostTensor* get(const std::string& key, std::vector<int64_t> shape, std::string* err) const;
  bool fold_bn(const std::string& bn, const std::vector<float>& bias, std::vector<float>* scale,
               std::vector<float>* shift, std::string* err) const;
  size_t elem_size() const { return dtype_ == 1 ? 2 : 4; }
  int chan_pad(int c) const { return dtype_ == 1 ? pad16(c) : c; }
  void release();

  std::vector<int> layer_sizes_;
  int start_idx_, end_idx_, num_classes_, dtype_, max_batch_;
  int device_ = 0, num_sms_ = 148;
  bool finalized_ = false;
  std::map<std::string, HostTensor> params_;
  std::vector<ConvOp> ops_;
  Dims in_dims_{}, out_dims_{};  // logical activation dims at the range boundaries
  int final_slot_ = SLOT_X;
  // bf16 mode: fold each downsample shortcut's temporal convolution into the block's last
  // convolution (RNB_B200_FUSE_SHORTCUT=0 keeps it a launch of its own + a residual read)
  bool fuse_shortcut_ = true;
  void* slots_[NUM_SLOTS] = {};
  size_t slot_elems_per_clip_[NUM_SLOTS] = {};
  void* d_classify_scratch_ = nullptr;  // staged logits + arrival counter of the fused classifier
  unsigned long long* d_dual_flags_ = nullptr;  // per-clip counters of the co-scheduled launches
  // bf16 handles that start at layer 1 run the stem as one fused kernel on a space-to-depth input
  // (conv_stem.cuh); RNB_B200_STEM=unfolded keeps the two patch-kernel launches of round 1
  bool stem_fused_ = false;
  void* d_stem_w_ = nullptr;  // spatial weights in the fused kernel's packing
  float* d_fc_w_ = nullptr;
  float* d_fc_b_ = nullptr;
  // One kernel launch of the tensor-core plan: convolution `op` over `batch` clips starting at clip
  // `first`: the patch kernel (patch), else the tap kernel, its CTA-pair variant (tc.pair) or its
  // split-K variant (tc.ks > 1).
  struct Launch {
    int op = 0;
    int first = 0, batch = 0;
    bool patch = false;
    ConvTcParams tc = {};
    ConvPatchParams pp = {};
    // Co-scheduled SpatioTemporalConv (conv_patch_dual_kernel): this (spatial, patch) launch also
    // runs the NEXT op, the temporal convolution reading its output; `merged` marks that next
    // entry, which is then not launched on its own.
    // Fused stem (conv_stem.cuh): this launch runs both convolutions of the stem; the next entry is
    // `merged`.
    bool stem_fused = false;
    StemFusedParams sf = {};
    // taps-along-N temporal kernel (conv_tsum.cuh) instead of the patch kernel
    bool tsum = false;
    ConvTsumParams ts = {};
    int dual_n_sp = 0;   // > 0: CTAs that start on the spatial convolution
    int dual_chunk = 0;  // temporal items per claim
    bool merged = false;
  };
  bool plan_stem(int batch, std::vector<Launch>* plan, std::string* err);
  bool dual_candidate(const std::vector<Launch>& plan, int i) const;
  void set_clip_order(std::vector<Launch>* plan) const;
  bool tune_dual(std::vector<Launch>* plan, int i, cudaStream_t stream);
  cudaError_t launch_one(const std::vector<Launch>& plan, size_t i, cudaStream_t stream) const;
  bool plan_conv(int op_index, int first, int batch, bool chunked, cudaStream_t stream,
                 const Launch* like, Launch* out, std::string* err, bool tune = true);
  bool build_plan(int batch, cudaStream_t stream, std::vector<Launch>* plan, std::string* err);
  bool derive_plan(int batch, std::vector<Launch>* plan, std::string* err);
  const std::vector<Launch>* ensure_plan(int batch, cudaStream_t stream, std::string* err,
                                         bool tune);
  uint64_t plan_key(int batch) const;
  bool load_cached_plan(const char* path, int batch, cudaStream_t stream, std::vector<Launch>* plan);
  void save_cached_plan(const char* path, int batch, const std::vector<Launch>& plan);
  cudaError_t launch_tc(const ConvTcParams& tc, cudaStream_t stream) const;
  std::map<int, std::vector<Launch>> plans_;  // per batch size
  std::set<int> tuned_;                       // batch sizes whose plan was timed (or read from the cache)
  int last_batch_ = 0;                        // batch size of the most recent forward
  int last_launches_ = 0;                     // kernels launched by the most recent forward
};

}  // namespace rnb
