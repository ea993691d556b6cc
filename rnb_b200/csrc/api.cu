// extern "C" surface of librnb_b200.so (declared in include/rnb_b200.h).
#include <atomic>
#include <string>
#include <vector>

#include "../../include/rnb_b200.h"
#include "kernels.cuh"
#include "model.cuh"

static_assert(RNB_STATIC_CLASSIFY_MAX_BATCH == rnb::kStaticClassifyMaxBatch &&
                  RNB_STATIC_CLASSIFY_MAX_CLASSES == rnb::kStaticClassifyMaxClasses,
              "include/rnb_b200.h and kernels.cuh disagree on the built-in scratch size");

namespace {

thread_local std::string g_last_error;

int fail(int code, const std::string& msg) {
  g_last_error = msg;
  return code;
}

int cuda_fail(cudaError_t e, const char* what) {
  (void)cudaGetLastError();
  return fail(RNB_ECUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

}  // namespace

extern "C" {

const char* rnb_last_error(void) { return g_last_error.c_str(); }

int rnb_init(int device) {
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceCount");
  if (device < 0 || device >= count)
    return fail(RNB_EINVAL, "device " + std::to_string(device) + " out of range (" +
                                std::to_string(count) + " visible)");
  e = cudaSetDevice(device);
  if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
  int major = 0;
  e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
  if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceGetAttribute");
  if (major != 10)
    return fail(RNB_EINVAL, "rnb_b200 is built for sm_100a only; device " + std::to_string(device) +
                                " has compute capability major " + std::to_string(major));
  return RNB_OK;
}

int rnb_r2p1d_create(const int* layer_sizes, int n_layer_sizes, int start_idx, int end_idx,
                     int num_classes, int dtype, int max_batch, void** out_handle) {
  if (!out_handle) return fail(RNB_EINVAL, "out_handle is NULL");
  *out_handle = nullptr;
  // same range check as R2P1DRunner.__init__ (models/r2p1d/model.py:38-44)
  if (start_idx < 1) return fail(RNB_EINVAL, "start_index must be >= 1");
  if (end_idx > 5) return fail(RNB_EINVAL, "end_index must be <= 5");
  if (start_idx > end_idx) return fail(RNB_EINVAL, "start_index must be <= end_index");
  if (dtype != RNB_DTYPE_F32 && dtype != RNB_DTYPE_BF16) return fail(RNB_EINVAL, "unknown dtype");
  if (max_batch < 1) return fail(RNB_EINVAL, "max_batch must be >= 1");
  if (num_classes < 1) return fail(RNB_EINVAL, "num_classes must be >= 1");
  std::vector<int> ls;
  if (layer_sizes) {
    if (n_layer_sizes != end_idx - start_idx + 1)
      return fail(RNB_EINVAL, "layer_sizes needs one entry per layer of the range (stem included)");
    ls.assign(layer_sizes, layer_sizes + n_layer_sizes);
  }
  *out_handle = new rnb::R2P1DModel(ls, start_idx, end_idx, num_classes, dtype, max_batch);
  return RNB_OK;
}

int rnb_r2p1d_load_param(void* handle, const char* key, const float* host_data, const int64_t* shape,
                         int ndim) {
  if (!handle || !key || !host_data || (ndim > 0 && !shape) || ndim < 0 || ndim > 8)
    return fail(RNB_EINVAL, "bad argument to rnb_r2p1d_load_param");
  std::string err;
  if (!static_cast<rnb::R2P1DModel*>(handle)->load_param(key, host_data, shape, ndim, &err))
    return fail(RNB_ESTATE, err);
  return RNB_OK;
}

int rnb_r2p1d_finalize(void* handle) {
  if (!handle) return fail(RNB_EINVAL, "handle is NULL");
  std::string err;
  const int rc = static_cast<rnb::R2P1DModel*>(handle)->finalize(&err);
  if (rc != RNB_OK) {
    (void)cudaGetLastError();
    return fail(rc, err);
  }
  return RNB_OK;
}

int rnb_r2p1d_forward(void* handle, const void* in, int in_fmt, int batch, void* out, int out_fmt,
                      void* stream) {
  if (!handle) return fail(RNB_EINVAL, "handle is NULL");
  std::string err;
  const int rc = static_cast<rnb::R2P1DModel*>(handle)->forward(
      in, in_fmt, batch, out, out_fmt, static_cast<cudaStream_t>(stream), &err);
  if (rc != RNB_OK) {
    (void)cudaGetLastError();
    return fail(rc, err);
  }
  return RNB_OK;
}

int rnb_r2p1d_prepare(void* handle, int batch, void* stream) {
  if (!handle) return fail(RNB_EINVAL, "handle is NULL");
  std::string err;
  const int rc = static_cast<rnb::R2P1DModel*>(handle)->prepare(
      batch, static_cast<cudaStream_t>(stream), &err);
  if (rc != RNB_OK) {
    (void)cudaGetLastError();
    return fail(rc, err);
  }
  return RNB_OK;
}

int rnb_r2p1d_force_kernel(void* handle, int batch, int op, int kind, int cm, int cn, int ks,
                           int block_n, int pair) {
  if (!handle) return fail(RNB_EINVAL, "handle is NULL");
  std::string err;
  const int rc = static_cast<rnb::R2P1DModel*>(handle)->force_kernel(batch, op, kind, cm, cn, ks,
                                                                      block_n, pair, &err);
  if (rc != RNB_OK) {
    (void)cudaGetLastError();
    return fail(rc, err);
  }
  return RNB_OK;
}

int rnb_r2p1d_plan_entry(void* handle, int batch, int op, int out7[7]) {
  auto* m = static_cast<rnb::R2P1DModel*>(handle);
  if (!m || !out7) return fail(RNB_EINVAL, "bad argument to rnb_r2p1d_plan_entry");
  if (!m->plan_entry(batch, op, out7))
    return fail(RNB_ESTATE, "no launch plan for this batch size and convolution yet");
  return RNB_OK;
}

int rnb_r2p1d_classify(void* handle, const void* in, int in_fmt, int batch, float* logits,
                       float* sum, int* cls, void* stream) {
  if (!handle) return fail(RNB_EINVAL, "handle is NULL");
  if (batch < 1) return fail(RNB_EINVAL, "rnb_r2p1d_classify needs at least one clip");
  std::string err;
  const int rc = static_cast<rnb::R2P1DModel*>(handle)->forward(
      in, in_fmt, batch, logits, RNB_FMT_F32_LOGITS, static_cast<cudaStream_t>(stream), &err,
      nullptr, sum, cls, true);
  if (rc != RNB_OK) {
    (void)cudaGetLastError();
    return fail(rc, err);
  }
  return RNB_OK;
}

size_t rnb_pool_fc_classify_scratch_bytes(int batch, int classes) {
  if (batch < 0 || classes < 0) return 0;
  return rnb::pool_fc_classify_scratch_bytes(batch, classes);
}

int rnb_pool_fc_classify(const void* x, int batch, int positions, int channels, const float* w,
                         const float* bias, int classes, float* logits, float* sum, int* cls,
                         void* scratch, size_t scratch_bytes, void* stream) {
  if (batch < 0 || positions < 1 || classes < 1 || channels < 8 || channels % 8 != 0)
    return fail(RNB_EINVAL, "rnb_pool_fc_classify: bad shape (channels must be a multiple of 8)");
  if (batch == 0) return RNB_OK;
  if (!x || !w || !bias || !scratch) return fail(RNB_EINVAL, "null pointer");
  if (scratch_bytes < rnb::pool_fc_classify_scratch_bytes(batch, classes))
    return fail(RNB_EINVAL, "rnb_pool_fc_classify: scratch buffer too small, need " +
                                std::to_string(rnb::pool_fc_classify_scratch_bytes(batch, classes)) +
                                " bytes");
  if (reinterpret_cast<uintptr_t>(scratch) % 256 != 0)
    return fail(RNB_EINVAL, "rnb_pool_fc_classify: scratch must be 256-byte aligned");
  cudaError_t e = rnb::launch_pool_fc_classify_bf16(
      static_cast<const __nv_bfloat16*>(x), batch, positions, channels, w, bias, classes, logits, sum,
      cls, scratch, static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return cuda_fail(e, "rnb_pool_fc_classify");
  return RNB_OK;
}

int rnb_pool_fc_classify_static(const void* x, int batch, int positions, int channels,
                                const float* w, const float* bias, int classes, float* logits,
                                float* sum, int* cls, void* stream) {
  if (batch < 0 || positions < 1 || classes < 1 || channels < 8 || channels % 8 != 0)
    return fail(RNB_EINVAL,
                "rnb_pool_fc_classify_static: bad shape (channels must be a multiple of 8)");
  if (batch > RNB_STATIC_CLASSIFY_MAX_BATCH || classes > RNB_STATIC_CLASSIFY_MAX_CLASSES)
    return fail(RNB_EINVAL,
                "rnb_pool_fc_classify_static: batch " + std::to_string(batch) + " x classes " +
                    std::to_string(classes) + " exceeds the built-in scratch (at most " +
                    std::to_string(RNB_STATIC_CLASSIFY_MAX_BATCH) + " clips x " +
                    std::to_string(RNB_STATIC_CLASSIFY_MAX_CLASSES) +
                    " classes); use rnb_pool_fc_classify with your own scratch buffer");
  if (batch == 0) return RNB_OK;
  if (!x || !w || !bias) return fail(RNB_EINVAL, "null pointer");
  cudaError_t e = rnb::launch_pool_fc_classify_bf16_static(
      static_cast<const __nv_bfloat16*>(x), batch, positions, channels, w, bias, classes, logits, sum,
      cls, static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return cuda_fail(e, "rnb_pool_fc_classify_static");
  return RNB_OK;
}

int rnb_r2p1d_profile(void* handle, const void* in, int in_fmt, int batch, void* out, int out_fmt,
                      void* stream, float* ms, int cap) {
  if (!handle) return fail(RNB_EINVAL, "handle is NULL");
  std::string err;
  const int rc = static_cast<rnb::R2P1DModel*>(handle)->profile(
      in, in_fmt, batch, out, out_fmt, static_cast<cudaStream_t>(stream), ms, cap, &err);
  if (rc != RNB_OK) {
    (void)cudaGetLastError();
    return fail(rc, err);
  }
  return RNB_OK;
}

int rnb_r2p1d_op_info(void* handle, int op, char* name, int name_cap, double* flops_per_clip,
                      double* act_bytes_per_clip, double* weight_bytes, int* kernel_kind) {
  auto* m = static_cast<rnb::R2P1DModel*>(handle);
  if (!m || !m->finalized()) return fail(RNB_ESTATE, "handle not finalized");
  std::string n;
  if (!m->op_info(op, &n, flops_per_clip, act_bytes_per_clip, weight_bytes, kernel_kind))
    return fail(RNB_EINVAL, "op index out of range");
  if (name && name_cap > 0) {
    snprintf(name, (size_t)name_cap, "%s", n.c_str());
  }
  return RNB_OK;
}

int rnb_r2p1d_io_shape(void* handle, int which, int64_t shape4[4]) {
  if (!handle || !shape4 || (which != 0 && which != 1))
    return fail(RNB_EINVAL, "bad argument to rnb_r2p1d_io_shape");
  static_cast<rnb::R2P1DModel*>(handle)->io_shape(which, shape4);
  return RNB_OK;
}

double rnb_r2p1d_flops(void* handle, int batch) {
  auto* m = static_cast<rnb::R2P1DModel*>(handle);
  if (!m || !m->finalized()) return 0.0;
  return m->flops(batch);
}

int rnb_r2p1d_num_launches(void* handle) {
  auto* m = static_cast<rnb::R2P1DModel*>(handle);
  if (!m || !m->finalized()) return fail(RNB_ESTATE, "handle not finalized");
  return m->num_launches();
}

int rnb_r2p1d_num_convs(void* handle) {
  auto* m = static_cast<rnb::R2P1DModel*>(handle);
  if (!m || !m->finalized()) return fail(RNB_ESTATE, "handle not finalized");
  return m->num_ops();
}

void rnb_r2p1d_destroy(void* handle) { delete static_cast<rnb::R2P1DModel*>(handle); }

int rnb_preprocess_u8(const uint8_t* in, int batch, const float* scale3, const float* bias3,
                      void* out, int out_fmt, void* stream) {
  if (batch < 0) return fail(RNB_EINVAL, "negative batch");
  if (batch == 0) return RNB_OK;
  if (!in || !out) return fail(RNB_EINVAL, "null pointer");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  cudaError_t e;
  if (out_fmt == RNB_FMT_F32_NCDHW) {
    e = rnb::launch_pre_u8_to_f32_ncdhw(in, batch, scale3, bias3, static_cast<float*>(out), s);
  } else if (out_fmt == RNB_FMT_BF16_NDHWC) {
    e = rnb::launch_pre_u8_to_bf16_ndhwc(in, batch, scale3, bias3,
                                         static_cast<__nv_bfloat16*>(out), 16, s);
  } else {
    return fail(RNB_EINVAL, "rnb_preprocess_u8: out_fmt must be F32_NCDHW or BF16_NDHWC");
  }
  if (e != cudaSuccess) return cuda_fail(e, "rnb_preprocess_u8");
  return RNB_OK;
}

int rnb_batch_gather(const void* const* srcs, const int* rows, int k, void* dst, size_t row_bytes,
                     void* stream) {
  if (k < 0 || k > rnb::kMaxGatherSrcs)
    return fail(RNB_EINVAL, "rnb_batch_gather: k must be in [0, " +
                                std::to_string(rnb::kMaxGatherSrcs) + "]");
  if (k == 0) return RNB_OK;
  if (!srcs || !rows || !dst) return fail(RNB_EINVAL, "null pointer");
  for (int i = 0; i < k; ++i)
    if (rows[i] < 0 || (rows[i] > 0 && !srcs[i]))
      return fail(RNB_EINVAL, "rnb_batch_gather: bad source " + std::to_string(i));
  cudaError_t e = rnb::launch_batch_gather(srcs, rows, k, dst, row_bytes,
                                           static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return cuda_fail(e, "rnb_batch_gather");
  return RNB_OK;
}

int rnb_logits_sum_argmax(const float* logits, int rows, int classes, float* sum, int* cls,
                          void* stream) {
  if (!logits || rows < 0 || classes < 1) return fail(RNB_EINVAL, "bad argument");
  cudaError_t e = rnb::launch_logits_sum_argmax(logits, rows, classes, sum, cls,
                                                static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return cuda_fail(e, "rnb_logits_sum_argmax");
  return RNB_OK;
}

namespace {
// Direct peer access between two devices, enabled once per ordered pair.  cudaMemcpyPeerAsync works
// without it (the driver then stages the copy through host memory); with it the copy engine moves
// the bytes over NVLink.  Returns cudaSuccess, or the error that says the pair is not reachable.
cudaError_t ensure_peer_access(int from_dev, int to_dev) {
  constexpr int kMax = 64;
  static std::atomic<int> state[kMax][kMax] = {};  // 0 unknown, 1 enabled, 2 not reachable
  if (from_dev < 0 || to_dev < 0 || from_dev >= kMax || to_dev >= kMax) return cudaErrorInvalidDevice;
  const int st = state[from_dev][to_dev].load();
  if (st == 1) return cudaSuccess;
  if (st == 2) return cudaErrorPeerAccessUnsupported;
  int can = 0;
  cudaError_t e = cudaDeviceCanAccessPeer(&can, from_dev, to_dev);
  if (e != cudaSuccess) return e;
  if (!can) {
    state[from_dev][to_dev].store(2);
    return cudaErrorPeerAccessUnsupported;
  }
  int cur = 0;
  e = cudaGetDevice(&cur);
  if (e != cudaSuccess) return e;
  if (cur != from_dev) e = cudaSetDevice(from_dev);
  if (e == cudaSuccess) {
    e = cudaDeviceEnablePeerAccess(to_dev, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) {
      (void)cudaGetLastError();
      e = cudaSuccess;
    }
  }
  if (cur != from_dev) (void)cudaSetDevice(cur);
  if (e == cudaSuccess) state[from_dev][to_dev].store(1);
  return e;
}

int peer_copy_impl(void* dst, int dst_dev, const void* src, int src_dev, size_t bytes,
                   cudaStream_t src_stream, bool fence_src, cudaStream_t copy_stream,
                   const char* what) {
  if (bytes == 0) return RNB_OK;
  if (!dst || !src) return fail(RNB_EINVAL, "null pointer");
  cudaError_t e = cudaSuccess;
  if (dst_dev != src_dev) {
    // both directions: the copy engine of either device may be the one that moves the bytes
    e = ensure_peer_access(dst_dev, src_dev);
    if (e == cudaSuccess) e = ensure_peer_access(src_dev, dst_dev);
    if (e != cudaSuccess) {
      (void)cudaGetLastError();
      return fail(RNB_ECUDA, std::string(what) + ": devices " + std::to_string(src_dev) + " and " +
                                 std::to_string(dst_dev) +
                                 " cannot reach each other's memory directly: " +
                                 cudaGetErrorString(e));
    }
  }
  if (fence_src && src_stream != copy_stream) {
    // order the copy after the work that produced `src`: an event on the producer's stream (created
    // on the producer's device) that the copy stream waits for on the GPU -- no host wait
    int cur = 0;
    e = cudaGetDevice(&cur);
    if (e != cudaSuccess) return cuda_fail(e, what);
    if (cur != src_dev) e = cudaSetDevice(src_dev);
    cudaEvent_t ev = nullptr;
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventRecord(ev, src_stream);
    if (cur != src_dev) (void)cudaSetDevice(cur);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(copy_stream, ev, 0);
    if (ev) (void)cudaEventDestroy(ev);  // released once the recorded work has completed
    if (e != cudaSuccess) return cuda_fail(e, what);
  }
  e = (dst_dev == src_dev) ? cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, copy_stream)
                           : cudaMemcpyPeerAsync(dst, dst_dev, src, src_dev, bytes, copy_stream);
  if (e != cudaSuccess) return cuda_fail(e, what);
  return RNB_OK;
}
}  // namespace

int rnb_peer_copy(void* dst, int dst_dev, const void* src, int src_dev, size_t bytes,
                  void* stream) {
  return peer_copy_impl(dst, dst_dev, src, src_dev, bytes, nullptr, false,
                        static_cast<cudaStream_t>(stream), "rnb_peer_copy");
}

int rnb_peer_copy_fenced(void* dst, int dst_dev, const void* src, int src_dev, size_t bytes,
                         void* src_stream, void* copy_stream) {
  return peer_copy_impl(dst, dst_dev, src, src_dev, bytes, static_cast<cudaStream_t>(src_stream),
                        true, static_cast<cudaStream_t>(copy_stream), "rnb_peer_copy_fenced");
}

}  // extern "C"
