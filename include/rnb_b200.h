/* rnb_b200 -- C ABI of the B200-native R(2+1)D hot path of the RnB video-inference benchmark.
 *
 * Every entry point takes plain pointers and sizes (no torch types), returns 0 on success or a
 * negative RNB_E* code, never throws, and is asynchronous with respect to the CUDA stream it is
 * given: the caller synchronises, exactly as RnB's runner does after calling a step
 * (reference runner.py:125-128).  `stream` is a cudaStream_t passed as void*; pass the stream the
 * runner made current (reference runner.py:41-45), never rely on the legacy default stream.
 * rnb_last_error() returns a human-readable message for the calling thread's last failure.
 *
 * Entry points that DO wait on the GPU, and say so below: rnb_r2p1d_finalize (uploads),
 * rnb_r2p1d_prepare (times candidate kernels) and rnb_r2p1d_profile.  Nothing else synchronises.
 *
 * Threads: a handle belongs to one host thread at a time; different threads may drive different
 * handles and devices concurrently (per-device one-time set-up inside the library is idempotent).
 *
 * Device-wide side effect: while at least one finalized bf16 handle exists on a device, the library
 * may set cudaLimitPersistingL2CacheSize on it (the convolution outputs are marked "persisting" so
 * that the next layer finds them in L2); destroying the last handle restores the previous limit.
 * RNB_B200_L2_PERSIST_MB=0 in the environment switches the feature off.
 *
 * File:line citations below are into the reference tree (snuspl/rnb @ 938d079) and name the
 * reference code each entry point replaces.
 */
#ifndef RNB_B200_H_
#define RNB_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RNB_OK 0
#define RNB_EINVAL (-1)   /* bad argument / unsupported configuration        */
#define RNB_ECUDA (-2)    /* a CUDA runtime or driver call failed             */
#define RNB_EPARAM (-3)   /* a state_dict entry is missing or has a bad shape */
#define RNB_ESTATE (-4)   /* call order violated (e.g. forward before finalize) */

/* arithmetic mode of a model handle */
#define RNB_DTYPE_F32 0   /* fp32 direct convolutions (parity mode)                     */
#define RNB_DTYPE_BF16 1  /* bf16 operands, fp32 accumulation on tcgen05 tensor cores   */

/* tensor formats at the boundary */
#define RNB_FMT_U8_NTCHW 0    /* uint8 (B, 8, 3, 112, 112): loader frames, models/r2p1d/model.py:140-147 */
#define RNB_FMT_F32_NCDHW 1   /* fp32 (B, C, T, H, W): RnB's inter-step tensors, control.py:41            */
#define RNB_FMT_BF16_NDHWC 2  /* bf16 (B, T, H, W, pad16(C)): this library's private hand-off format      */
#define RNB_FMT_F32_LOGITS 3  /* fp32 (B, num_classes): only as the output of a model ending at layer 5  */

/* Select `device` for the calling thread and check that it is an sm_100 GPU.
 * Replaces torch.cuda.set_device(g_idx) at runner.py:31-32 for code that does not go through torch. */
int rnb_init(int device);

/* Build layers start_idx..end_idx (1 = stem ... 5 = conv5 + pool + classifier) of R(2+1)D on the
 * current device.  `layer_sizes` has one entry per layer of the range, stem position included and
 * ignored -- the indexing rule of R2Plus1DLayerNet (models/r2p1d/network.py:18-34) as driven by
 * R2P1DRunner (models/r2p1d/model.py:35-49); NULL means two blocks per stage (model.py:46-47).
 * Replaces R2Plus1DLayerWrapper(...).to(device) at models/r2p1d/model.py:49 and
 * R2Plus1DClassifier(...) at model.py:177. */
int rnb_r2p1d_create(const int* layer_sizes, int n_layer_sizes, int start_idx, int end_idx,
                     int num_classes, int dtype, int max_batch, void** out_handle);

/* Hand over one state_dict entry (host fp32, torch layout) under its reference key, e.g.
 * "res2plus1d.conv2.block1.conv1.spatial_conv.weight" or "linear.bias".  Keys outside the layer
 * range and "num_batches_tracked" entries are accepted and ignored, so the prefix filter of
 * models/r2p1d/model.py:52-63 is not needed.  Replaces model.load_state_dict at model.py:63,182. */
int rnb_r2p1d_load_param(void* handle, const char* key, const float* host_data, const int64_t* shape,
                         int ndim);

/* (No counterpart in the reference, where torch modules hold the parameters.)  Fold BatchNorm
 * (inference statistics) and biases into per-channel scale/shift, repack the
 * weights for the tensor cores, upload everything and allocate the activation arena.  After this
 * call the handle allocates no device memory again (a batch size it has not seen costs a few
 * kilobytes of host memory for its launch plan). */
int rnb_r2p1d_finalize(void* handle);

/* Choose the kernels for `batch` clips by timing the candidates of every layer on the GPU -- the
 * counterpart of torch.backends.cudnn.benchmark = True at runner.py:25, done explicitly and ahead of
 * the serving loop (R2P1DRunner's warm-up forwards, models/r2p1d/model.py:66-71, are where the
 * reference pays for it).  Synchronises `stream` and overwrites the handle's activation arena; call
 * it from the step constructor for every batch size the pipeline can produce, never inside stream
 * capture.  RNB_B200_PLAN_CACHE=<file> stores / reuses the choices so that processes agree.
 * rnb_r2p1d_forward on a batch size that was never prepared does NOT time anything: it derives a
 * plan on the host from the nearest prepared size (or from fixed rules) and stays asynchronous. */
int rnb_r2p1d_prepare(void* handle, int batch, void* stream);

/* Test and sweep hook (no counterpart in the reference): pin the kernel that runs convolution `op`
 * (0-based, as in rnb_r2p1d_op_info) when the handle is called with `batch` clips.  kind: 1 =
 * conv_tc_kernel with a cm x cn thread-block cluster and channel tile block_n (0 = default), 2 =
 * conv_patch_kernel (pair = 1: its CTA-pair form), 3 = conv_splitk_kernel with ks CTAs per tile and
 * channel tile <= block_n, 4 = conv_pair_kernel with channel tile <= block_n, 5 = conv_pair_kernel
 * in temporal-patch mode, 6 = conv_patch_dual_kernel: ONE launch for the 1 x k x k convolution `op`
 * and the k x 1 x 1 convolution op + 1 of the same SpatioTemporalConv, with cm CTAs starting on the
 * spatial half (0: three quarters of the SMs) and cn temporal tiles per claim (0: 4), 8 =
 * conv_tsum_kernel (3 x 1 x 1 layers of 8-frame clips with 64 output channels).  RNB_EINVAL if
 * that kernel or shape does not apply to the layer. */
int rnb_r2p1d_force_kernel(void* handle, int batch, int op, int kind, int cm, int cn, int ks,
                           int block_n, int pair);

/* The other half of the hook (no counterpart in the reference): what the plan of `batch` clips
 * currently runs convolution `op` with -- out7 = {kind, cm, cn, ks, block_n, pair, tuned}, the
 * arguments rnb_r2p1d_force_kernel takes plus 1 if the plan was timed by rnb_r2p1d_prepare.
 * RNB_ESTATE while the handle has no plan for that batch size. */
int rnb_r2p1d_plan_entry(void* handle, int batch, int op, int out7[7]);

/* One forward pass over `batch` clips.  `in`/`out` are device pointers.
 *   in_fmt : U8_NTCHW (only when start_idx == 1; fuses the loader's float()+permute of
 *            models/r2p1d/model.py:149-150), F32_NCDHW (shape of model.py:29-33) or BF16_NDHWC.
 *   out_fmt: F32_LOGITS when end_idx == 5, else F32_NCDHW or BF16_NDHWC.
 * Replaces self.model(tensor) at models/r2p1d/model.py:84 and model.py:228. */
int rnb_r2p1d_forward(void* handle, const void* in, int in_fmt, int batch, void* out, int out_fmt,
                      void* stream);

/* Forward pass plus the per-video reduction of the aggregator step, for a handle ending at layer 5:
 * sum[j] = sum over the `batch` clips of logits[b][j] and cls = argmax_j sum[j] (lowest index on
 * ties), produced on the GPU by the classifier kernel itself.  logits (batch x num_classes) and sum
 * (num_classes) may be NULL.  Replaces R2P1DAggregator's tensor.cpu().numpy().sum(axis=0) and
 * .argmax() (models/r2p1d/model.py:254,259) when runner and aggregator share a GPU, so that one
 * int instead of (n, 400) floats crosses to the host. */
int rnb_r2p1d_classify(void* handle, const void* in, int in_fmt, int batch, float* logits,
                       float* sum, int* cls, void* stream);

/* Per-clip logical shape (C, T, H, W) of the handle's input (which = 0) or output (which = 1);
 * for a model ending at layer 5 the output is (num_classes, 1, 1, 1).  Mirrors
 * R2P1DRunner.input_dict at models/r2p1d/model.py:29-33. */
int rnb_r2p1d_io_shape(void* handle, int which, int64_t shape4[4]);

/* Bookkeeping for benchmarks (no counterpart in the reference): algorithmic FLOPs of one forward
 * over `batch` clips (2 x MACs of every convolution and the classifier), the number of kernels the
 * most recent forward launched (a convolution may run in several batch chunks, RNB_B200_CHUNK_MB),
 * and the number of convolutions (Conv3d modules) in the layer range. */
double rnb_r2p1d_flops(void* handle, int batch);
int rnb_r2p1d_num_launches(void* handle);
int rnb_r2p1d_num_convs(void* handle);

/* Diagnostics for bench.py and profiles/ (no counterpart in the reference, which times steps with
 * wall-clock TimeCards only): the same forward pass with a CUDA event recorded on
 * `stream` between consecutive kernels.  With n = rnb_r2p1d_num_convs(): ms[0] = input staging,
 * ms[1 .. n] = the convolutions in module order (launches of one convolution summed), ms[n + 1] =
 * pool + classifier (or the output layout change); `cap` is the capacity of ms and must be
 * >= n + 2.  Synchronises `stream` before returning.
 * rnb_r2p1d_op_info() describes convolution `op` (0-based): its state_dict prefix, algorithmic
 * FLOPs per clip, and the bytes per clip it must move at least once (input + output + residual
 * activations in the handle's dtype; weights are per batch and reported separately), and which
 * kernel the launch plan of the most recent batch size runs it with (before the first forward: the
 * kernel that applies): 0 = fp32 direct convolution, 1 = conv_tc_kernel, 2 = conv_patch_kernel,
 * 3 = conv_splitk_kernel, 4 = conv_pair_kernel, 5 = conv_pair_kernel in temporal-patch mode, 6 =
 * conv_patch_dual_kernel, 7 = conv_stem_fused_kernel (both reported for the two convolutions of the
 * module they run in one launch; the second one's entry in rnb_r2p1d_profile is then 0), 8 =
 * conv_tsum_kernel (3x1x1 layers of 8-frame clips, filter taps stacked along the output columns). */
int rnb_r2p1d_profile(void* handle, const void* in, int in_fmt, int batch, void* out, int out_fmt,
                      void* stream, float* ms, int cap);
int rnb_r2p1d_op_info(void* handle, int op, char* name, int name_cap, double* flops_per_clip,
                      double* act_bytes_per_clip, double* weight_bytes, int* kernel_kind);

void rnb_r2p1d_destroy(void* handle);

/* uint8 loader frames (B, 8, 3, 112, 112) -> value * scale[c] + bias[c] in `out_fmt`
 * (F32_NCDHW = exactly frames.float().permute(0, 2, 1, 3, 4) made contiguous, which the reference
 * does at models/r2p1d/model.py:149-150 plus the strided copy at runner.py:167-170; or
 * BF16_NDHWC with 16 channel slots).  scale3/bias3 are HOST arrays of 3 floats or NULL (identity,
 * the reference behaviour). */
int rnb_preprocess_u8(const uint8_t* in, int batch, const float* scale3, const float* bias3,
                      void* out, int out_fmt, void* stream);

/* The classifier tail of the network as one launch: global average pool over `positions` (the
 * AdaptiveAvgPool3d(1) of models/r2p1d/network.py:35), Linear(channels -> classes)
 * (network.py:56,60), and the aggregator's sum over clips + arg-max (models/r2p1d/model.py:254,259).
 * x: bf16 (batch, positions, channels) channels-last activations of the last stage; w: fp32
 * (classes, channels); bias: fp32 (classes).  logits / sum / cls may each be NULL.
 * `scratch` is device memory of at least rnb_pool_fc_classify_scratch_bytes(batch, classes) bytes,
 * 256-byte aligned, owned by the caller; its contents on entry do not matter and two calls that may
 * overlap on the GPU must not share it. */
size_t rnb_pool_fc_classify_scratch_bytes(int batch, int classes);
int rnb_pool_fc_classify(const void* x, int batch, int positions, int channels, const float* w,
                         const float* bias, int classes, float* logits, float* sum, int* cls,
                         void* scratch, size_t scratch_bytes, void* stream);

/* rnb_pool_fc_classify for callers that cannot allocate device memory.  (No counterpart in the
 * reference.)
 *
 * Who needs it: code that runs where cudaMalloc is off limits -- a serving process whose device
 * memory was carved up once at start-up and then locked, a latency-critical loop that may not take
 * the allocator's lock or sit through cudaMalloc's implicit device synchronisation, or a host
 * application that only lends this library raw pointers and owns every allocation itself.  Such a
 * caller has nowhere to get the scratch buffer rnb_pool_fc_classify asks for, so this variant
 * brings its own: a fixed array compiled into the library, 256 bytes +
 * RNB_STATIC_CLASSIFY_MAX_BATCH x RNB_STATIC_CLASSIFY_MAX_CLASSES floats (about 256 KiB) per GPU
 * that loads the library, held until the process exits.  Same arguments and results as
 * rnb_pool_fc_classify minus the scratch pair; x, w, bias and the outputs are still the caller's.
 *
 *  - Calls no cudaMalloc / cudaFree or any other allocating routine, on any call.
 *  - batch > RNB_STATIC_CLASSIFY_MAX_BATCH or classes > RNB_STATIC_CLASSIFY_MAX_CLASSES is refused
 *    with RNB_EINVAL and a message that names the limits.
 *  - The first call on each device looks up the array's device address (with that device current)
 *    and remembers it; make that first call before you start capturing a stream.
 *  - All calls on one GPU share the one array.  Each call re-zeroes the part that must start at zero
 *    (the arrival counter), so stale contents from the previous call are harmless, but two calls
 *    that can overlap on the GPU (different streams, or different threads/processes' streams on the
 *    same device within one process) race on it: serialise them, or use rnb_pool_fc_classify with
 *    one scratch buffer per stream. */
#define RNB_STATIC_CLASSIFY_MAX_BATCH 64
#define RNB_STATIC_CLASSIFY_MAX_CLASSES 1024
int rnb_pool_fc_classify_static(const void* x, int batch, int positions, int channels,
                                const float* w, const float* bias, int classes, float* logits,
                                float* sum, int* cls, void* stream);

/* Concatenate k row-blocks (rows[i] rows of row_bytes bytes each, device pointers in the HOST
 * array srcs) into dst.  Replaces torch.cat(self.stacked_tensors, dim=0) at batcher.py:28. */
int rnb_batch_gather(const void* const* srcs, const int* rows, int k, void* dst, size_t row_bytes,
                     void* stream);

/* sum[j] = sum over rows of logits[r][j] (optional, may be NULL) and cls = argmax_j sum[j] with
 * numpy's rules: lowest index on ties, and a NaN sum counts as the maximum (the first NaN wins).
 * The same rules hold for rnb_r2p1d_classify and rnb_pool_fc_classify.  Replaces
 * tensor.cpu().numpy().sum(axis=0) / .argmax() at models/r2p1d/model.py:254,259,278. */
int rnb_logits_sum_argmax(const float* logits, int rows, int classes, float* sum, int* cls,
                          void* stream);

/* Copy `bytes` from src on device src_dev to dst on device dst_dev, ordered on `stream` ONLY: the
 * caller guarantees that whatever produced `src` has completed or is ordered before `stream`
 * (RnB's runner synchronises its stream before it publishes, runner.py:127-128).  When the devices
 * differ, direct peer access is enabled once per pair so that the bytes travel over NVLink; pairs
 * that cannot reach each other fail with RNB_ECUDA.  Replaces
 * placeholder[:n].copy_(shared[:n]) at runner.py:112-113 and shared[:n].copy_(segment) at
 * runner.py:170. */
int rnb_peer_copy(void* dst, int dst_dev, const void* src, int src_dev, size_t bytes, void* stream);

/* Same copy, also ordered after everything queued so far on `src_stream` (a stream of src_dev, the
 * producer's): an event recorded there is waited for by `copy_stream` on the GPU, the host never
 * blocks.  This is the ordering Tensor.copy_ gives the reference at runner.py:112-113,170 (it
 * synchronises both devices' current streams) without the host-side wait. */
int rnb_peer_copy_fenced(void* dst, int dst_dev, const void* src, int src_dev, size_t bytes,
                         void* src_stream, void* copy_stream);

const char* rnb_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* RNB_B200_H_ */
