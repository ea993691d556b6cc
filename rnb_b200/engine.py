"""Torch-facing wrapper of one ``rnb_r2p1d_*`` model handle.

PyTorch is plumbing here (device memory, streams); every FLOP runs in ``librnb_b200.so``.
"""
import ctypes
import math

import torch

from . import _lib

NUM_CLASSES = 400
# per-clip input shape by start layer and output shape by end layer
# (reference models/r2p1d/model.py:29-33; the output of layer i is the input of layer i+1)
LAYER_INPUT_SHAPE = {1: (3, 8, 112, 112), 2: (64, 8, 56, 56), 3: (64, 8, 56, 56),
                     4: (128, 4, 28, 28), 5: (256, 2, 14, 14)}
LAYER_OUTPUT_SHAPE = {1: (64, 8, 56, 56), 2: (64, 8, 56, 56), 3: (128, 4, 28, 28),
                      4: (256, 2, 14, 14)}
_STAGE = {2: (64, 64, False), 3: (64, 128, True), 4: (128, 256, True), 5: (256, 512, True)}


def _mid(cin, cout, kt, kh, kw):
    return int(math.floor((kt * kh * kw * cin * cout) / (kh * kw * cin + kt * cout)))


def _st_conv_entries(prefix, cin, cout, k):
    kt, kh, kw = k
    mid = _mid(cin, cout, kt, kh, kw)
    return [
        (prefix + '.spatial_conv.weight', (mid, cin, 1, kh, kw)),
        (prefix + '.spatial_conv.bias', (mid,)),
        (prefix + '.bn.weight', (mid,)), (prefix + '.bn.bias', (mid,)),
        (prefix + '.bn.running_mean', (mid,)), (prefix + '.bn.running_var', (mid,)),
        (prefix + '.temporal_conv.weight', (cout, mid, kt, 1, 1)),
        (prefix + '.temporal_conv.bias', (cout,)),
    ]


def _bn_entries(prefix, n):
    return [(prefix + '.weight', (n,)), (prefix + '.bias', (n,)),
            (prefix + '.running_mean', (n,)), (prefix + '.running_var', (n,))]


def state_dict_spec(start_index=1, end_index=5, num_classes=NUM_CLASSES, layer_sizes=None):
    """(key, shape) of every state_dict entry layers start..end need, in module order.

    Key names are the reference's (``res2plus1d.conv{i}...``, ``linear``; reference
    models/r2p1d/model.py:55-62), so a checkpoint trained on R2Plus1DClassifier loads unchanged.
    """
    if layer_sizes is None:
        layer_sizes = [2] * (end_index - start_index + 1)
    spec = []
    for pos, layer in enumerate(range(start_index, end_index + 1)):
        lp = 'res2plus1d.conv%d' % layer
        if layer == 1:
            spec += _st_conv_entries(lp, 3, 64, (3, 7, 7))
            continue
        cin, cout, down = _STAGE[layer]
        for blk in range(layer_sizes[pos]):
            bp = lp + ('.block1' if blk == 0 else '.blocks.%d' % (blk - 1))
            ds = down and blk == 0
            bcin = cin if blk == 0 else cout
            if ds:
                spec += _st_conv_entries(bp + '.downsampleconv', bcin, cout, (1, 1, 1))
                spec += _bn_entries(bp + '.downsamplebn', cout)
            spec += _st_conv_entries(bp + '.conv1', bcin, cout, (3, 3, 3))
            spec += _bn_entries(bp + '.bn1', cout)
            spec += _st_conv_entries(bp + '.conv2', cout, cout, (3, 3, 3))
            spec += _bn_entries(bp + '.bn2', cout)
    if end_index == 5:
        spec += [('linear.weight', (num_classes, 512)), ('linear.bias', (num_classes,))]
    return spec


def random_state_dict(start_index=1, end_index=5, num_classes=NUM_CLASSES, layer_sizes=None,
                      seed=0):
    """Random-init weights with torch's default fan-in scaling (uniform +-1/sqrt(fan_in)).

    Stands in for the reference's checkpoint (models/r2p1d/model.py:18,50), which is not
    distributed; BatchNorm entries are identity statistics.
    """
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for key, shape in state_dict_spec(start_index, end_index, num_classes, layer_sizes):
        leaf = key.rsplit('.', 1)[1]
        parent = key.rsplit('.', 2)[-2]
        if parent.startswith('bn') or parent.endswith('bn'):
            if leaf in ('weight', 'running_var'):
                sd[key] = torch.ones(shape)
            else:
                sd[key] = torch.zeros(shape)
        elif leaf == 'weight':
            fan_in = 1
            for d in shape[1:]:
                fan_in *= d
            bound = 1.0 / math.sqrt(fan_in)
            sd[key] = (torch.rand(shape, generator=g) * 2 - 1) * bound
            sd['_fan_in:' + key.rsplit('.', 1)[0]] = fan_in
        else:
            fan_in = sd.pop('_fan_in:' + key.rsplit('.', 1)[0])
            bound = 1.0 / math.sqrt(fan_in)
            sd[key] = (torch.rand(shape, generator=g) * 2 - 1) * bound
    return sd


class R2P1DEngine:
    """Layers ``start_index..end_index`` of R(2+1)D on one GPU.

    ``dtype`` is ``'bf16'`` (tensor cores, the fast path) or ``'fp32'`` (parity mode).
    ``layer_sizes`` follows R2P1DRunner: one entry per layer of the range, stem position included
    (reference models/r2p1d/model.py:46-47, network.py:20-34).
    """

    def __init__(self, device, state_dict, start_index=1, end_index=5, num_classes=NUM_CLASSES,
                 layer_sizes=None, dtype='bf16', max_batch=64):
        self.lib = _lib.load()
        self.device = torch.device(device)
        if self.device.type != 'cuda':
            raise RuntimeError('rnb_b200 runs on CUDA devices only (got %s)' % (self.device,))
        self.start_index, self.end_index = start_index, end_index
        self.num_classes = num_classes
        self.dtype = dtype
        self.max_batch = max_batch
        dt = {'bf16': _lib.DTYPE_BF16, 'fp32': _lib.DTYPE_F32}[dtype]
        index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device_index = index
        self._handle = ctypes.c_void_p()
        with torch.cuda.device(index):
            _lib.check(self.lib.rnb_init(index))
            if layer_sizes is not None:
                arr = (ctypes.c_int * len(layer_sizes))(*[int(v) for v in layer_sizes])
                n = len(layer_sizes)
            else:
                arr, n = None, 0
            _lib.check(self.lib.rnb_r2p1d_create(arr, n, start_index, end_index, num_classes, dt,
                                                 max_batch, ctypes.byref(self._handle)))
            try:
                for key, value in state_dict.items():
                    if not torch.is_tensor(value) or key.endswith('num_batches_tracked'):
                        continue
                    host = value.detach().to('cpu', torch.float32).contiguous()
                    shape = (ctypes.c_int64 * max(host.dim(), 1))(*host.shape)
                    _lib.check(self.lib.rnb_r2p1d_load_param(
                        self._handle, key.encode(), ctypes.c_void_p(host.data_ptr()), shape,
                        host.dim()))
                _lib.check(self.lib.rnb_r2p1d_finalize(self._handle))
            except Exception:
                self.close()
                raise
        if end_index == 5:
            self.out_shape = (num_classes,)
        else:
            self.out_shape = LAYER_OUTPUT_SHAPE[end_index]
        self.in_shape = LAYER_INPUT_SHAPE[start_index]
        self._out = torch.empty((max_batch,) + self.out_shape, dtype=torch.float32,
                                device=self.device)

    # ------------------------------------------------------------------ lifetime
    def close(self):
        if getattr(self, '_handle', None) is not None and self._handle.value:
            self.lib.rnb_r2p1d_destroy(self._handle)
            self._handle = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001 - interpreter shutdown
            pass

    # ------------------------------------------------------------------ bookkeeping
    def flops(self, batch):
        return float(self.lib.rnb_r2p1d_flops(self._handle, batch))

    @property
    def num_launches(self):
        """Kernels the most recent forward launched."""
        return int(self.lib.rnb_r2p1d_num_launches(self._handle))

    @property
    def num_convs(self):
        return int(self.lib.rnb_r2p1d_num_convs(self._handle))

    def prepare(self, *batches):
        """Time the candidate kernels for each of ``batches`` clips and keep the fastest per layer.

        The explicit counterpart of ``cudnn.benchmark`` (reference runner.py:25): synchronises the
        current stream and overwrites the engine's activation arena, so call it during set-up.  A
        batch size that was never prepared still runs -- with a plan derived on the host from the
        nearest prepared size -- and ``forward`` stays asynchronous either way.
        """
        stream = torch.cuda.current_stream(self.device).cuda_stream
        with torch.cuda.device(self.device_index):
            for b in batches:
                if 1 <= int(b) <= self.max_batch:
                    _lib.check(self.lib.rnb_r2p1d_prepare(self._handle, int(b),
                                                          ctypes.c_void_p(stream)))
        return self

    KERNEL_KINDS = ('conv_ref', 'conv_tc', 'conv_patch', 'conv_splitk', 'conv_pair', 'conv_tpatch',
                    'conv_patch_dual', 'conv_stem_fused', 'conv_tsum')

    def force_kernel(self, batch, op, kind, cm=0, cn=0, ks=0, block_n=0, pair=0):
        """Pin the kernel of convolution ``op`` at ``batch`` clips (tests and sweeps).  ``kind`` is a
        name from ``KERNEL_KINDS``; raises RnbError when it does not apply to the layer."""
        k = self.KERNEL_KINDS.index(kind)
        _lib.check(self.lib.rnb_r2p1d_force_kernel(self._handle, int(batch), int(op), k, cm, cn, ks,
                                                   block_n, pair))

    def plan_entry(self, batch, op):
        """(kind, dict of force_kernel arguments, tuned?) of convolution ``op`` at ``batch`` clips."""
        out = (ctypes.c_int * 7)()
        _lib.check(self.lib.rnb_r2p1d_plan_entry(self._handle, int(batch), int(op), out))
        return (self.KERNEL_KINDS[out[0]],
                dict(cm=out[1], cn=out[2], ks=out[3], block_n=out[4], pair=out[5]), bool(out[6]))

    def op_info(self):
        """[(state_dict prefix, FLOPs per clip, activation bytes per clip, weight bytes, kernel)]
        per convolution; kernel is one of ``KERNEL_KINDS`` (what the launch plan of the most recent
        batch size runs)."""
        info = []
        for i in range(self.num_convs):
            name = ctypes.create_string_buffer(256)
            f, a, w, k = ctypes.c_double(), ctypes.c_double(), ctypes.c_double(), ctypes.c_int()
            _lib.check(self.lib.rnb_r2p1d_op_info(self._handle, i, name, 256, ctypes.byref(f),
                                                  ctypes.byref(a), ctypes.byref(w),
                                                  ctypes.byref(k)))
            info.append((name.value.decode(), f.value, a.value, w.value,
                         self.KERNEL_KINDS[k.value]))
        return info

    def profile(self, x):
        """Per-kernel milliseconds of one forward (CUDA events between launches); synchronises."""
        x, fmt, batch = self._check_input(x)
        out = self._out[:batch]
        out_fmt = _lib.FMT_F32_LOGITS if self.end_index == 5 else _lib.FMT_F32_NCDHW
        n = self.num_convs + 2
        ms = (ctypes.c_float * n)()
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(self.lib.rnb_r2p1d_profile(self._handle, ctypes.c_void_p(x.data_ptr()), fmt,
                                              batch, ctypes.c_void_p(out.data_ptr()), out_fmt,
                                              ctypes.c_void_p(stream), ms, n))
        return list(ms)

    # ------------------------------------------------------------------ forward
    def _check_input(self, x):
        batch = int(x.shape[0])
        if x.device != self.device and not (x.is_cuda and x.device.index == self.device_index):
            raise ValueError('input lives on %s, engine on %s' % (x.device, self.device))
        if x.dtype == torch.uint8:
            if tuple(x.shape[1:]) != (8, 3, 112, 112):
                raise ValueError('uint8 frames must be (B, 8, 3, 112, 112), got %s' % (tuple(x.shape),))
            fmt = _lib.FMT_U8_NTCHW
        elif x.dtype == torch.float32:
            if tuple(x.shape[1:]) != self.in_shape:
                raise ValueError('expected input (B, %s), got %s' % (self.in_shape, tuple(x.shape)))
            fmt = _lib.FMT_F32_NCDHW
        elif x.dtype == torch.bfloat16 and self.dtype == 'bf16' and self.start_index > 1:
            # the library's private hand-off format between layer ranges: channels last, 16-padded
            if tuple(x.shape[1:]) != self.wire_shape(self.in_shape):
                raise ValueError('expected bf16 input (B, %s), got %s'
                                 % (self.wire_shape(self.in_shape), tuple(x.shape)))
            fmt = _lib.FMT_BF16_NDHWC
        else:
            raise TypeError('input must be uint8 frames, float32 (B, C, T, H, W) activations or '
                            'bf16 channels-last activations, got %s' % x.dtype)
        if not x.is_contiguous():
            x = x.contiguous()
        if batch > self.max_batch:
            raise ValueError('batch %d exceeds max_batch %d' % (batch, self.max_batch))
        return x, fmt, batch

    @staticmethod
    def wire_shape(cthw):
        """Per-clip shape of the private bf16 hand-off format for a logical (C, T, H, W)."""
        c, t, h, w = cthw
        return (t, h, w, (c + 15) // 16 * 16)

    def forward(self, x, out=None, wire=False):
        """``x``: uint8 (B, 8, 3, 112, 112) loader frames (start layer 1 only), fp32
        (B, C, T, H, W), or -- between two engines -- bf16 (B, T, H, W, pad16(C)).  Returns fp32
        (B, num_classes) logits, or fp32 (B, C, T, H, W) activations when the range ends before
        layer 5 (``wire=True``: the bf16 channels-last form instead, half the bytes and no layout
        change on either side).  The result aliases an engine-owned buffer that the next call
        overwrites, unless ``out`` is given.  Asynchronous on the current stream.
        """
        x, fmt, batch = self._check_input(x)
        if wire:
            if self.end_index == 5 or self.dtype != 'bf16':
                raise ValueError('wire=True needs a bf16 engine that ends before layer 5')
            if out is None:
                if not hasattr(self, '_out_wire'):
                    self._out_wire = torch.empty((self.max_batch,) + self.wire_shape(self.out_shape),
                                                 dtype=torch.bfloat16, device=self.device)
                out = self._out_wire[:batch]
            stream = torch.cuda.current_stream(self.device).cuda_stream
            _lib.check(self.lib.rnb_r2p1d_forward(self._handle, ctypes.c_void_p(x.data_ptr()), fmt,
                                                  batch, ctypes.c_void_p(out.data_ptr()),
                                                  _lib.FMT_BF16_NDHWC, ctypes.c_void_p(stream)))
            return out
        if out is None:
            out = self._out[:batch]
        out_fmt = _lib.FMT_F32_LOGITS if self.end_index == 5 else _lib.FMT_F32_NCDHW
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(self.lib.rnb_r2p1d_forward(self._handle, ctypes.c_void_p(x.data_ptr()), fmt,
                                              batch, ctypes.c_void_p(out.data_ptr()), out_fmt,
                                              ctypes.c_void_p(stream)))
        return out

    __call__ = forward

    def classify(self, x):
        """Forward pass plus the per-video reduction on the GPU.

        Returns ``(logits, total, cls)``: fp32 (B, num_classes), fp32 (num_classes,) = sum over
        the clips, int32 scalar = arg-max of the sum -- engine-owned device tensors, asynchronous
        on the current stream (what R2P1DAggregator computes on the host, model.py:254,259).
        """
        if self.end_index != 5:
            raise ValueError('classify() needs a range that ends at layer 5')
        x, fmt, batch = self._check_input(x)
        if batch < 1:
            raise ValueError('classify() needs at least one clip')
        if not hasattr(self, '_sum'):
            self._sum = torch.empty(self.num_classes, dtype=torch.float32, device=self.device)
            self._cls = torch.empty((), dtype=torch.int32, device=self.device)
        out = self._out[:batch]
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(self.lib.rnb_r2p1d_classify(
            self._handle, ctypes.c_void_p(x.data_ptr()), fmt, batch,
            ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(self._sum.data_ptr()),
            ctypes.c_void_p(self._cls.data_ptr()), ctypes.c_void_p(stream)))
        return out, self._sum, self._cls


class HostPipeline:
    """Feeds an engine from pinned host memory, end to end: host uint8 frames in, host logits out.

    The host->device copy of batch i+1 runs on a copy stream while batch i is in the network
    (double-buffered device inputs, CUDA events in both directions); the logits of batch i land in
    a per-slot device buffer and go back to pinned host memory on a third stream, so the compute
    stream runs kernel to kernel (a copy between two forwards would break the programmatic
    dependent launch of the next staging kernel).  This is the serving loop a caller without
    device-resident data uses -- RnB's loader hands device tensors to the runner, so there the
    copy does not exist.
    """

    def __init__(self, engine, batch, depth=2):
        self.engine = engine
        self.depth = depth
        dev = engine.device
        self.copy_stream = torch.cuda.Stream(device=dev)
        self.out_stream = torch.cuda.Stream(device=dev)
        self.dev_in = [torch.empty((batch, 8, 3, 112, 112), dtype=torch.uint8, device=dev)
                       for _ in range(depth)]
        self.dev_out = [torch.empty((batch,) + engine.out_shape, dtype=torch.float32, device=dev)
                        for _ in range(depth)]
        self.fwd_done = [torch.cuda.Event() for _ in range(depth)]
        self.host_out = [torch.empty((batch,) + engine.out_shape, dtype=torch.float32).pin_memory()
                         for _ in range(depth)]
        self.in_ready = [torch.cuda.Event() for _ in range(depth)]
        self.in_free = [torch.cuda.Event() for _ in range(depth)]
        self.out_ready = [torch.cuda.Event() for _ in range(depth)]
        self.count = 0
        self.unread = [False] * depth  # slot holds a result the caller has not asked for yet
        self.h2d_bytes = self.dev_in[0].numel()
        self.d2h_bytes = self.host_out[0].numel() * 4

    def submit(self, host_frames):
        """Queue one batch (pinned uint8 (B, 8, 3, 112, 112)); returns the slot its logits land in.

        ``result(slot)`` waits for them.  At most ``depth`` batches may be outstanding.
        """
        slot = self.count % self.depth
        if self.unread[slot]:
            raise RuntimeError('HostPipeline: %d batches are outstanding; call result(%d) before '
                               'submitting another one (its host buffer would be overwritten)'
                               % (self.depth, slot))
        compute = torch.cuda.current_stream(self.engine.device)
        n = host_frames.shape[0]
        with torch.cuda.stream(self.copy_stream):
            if self.count >= self.depth:
                self.copy_stream.wait_event(self.in_free[slot])
            self.dev_in[slot][:n].copy_(host_frames, non_blocking=True)
            self.in_ready[slot].record(self.copy_stream)
        compute.wait_event(self.in_ready[slot])
        # (slot reuse: result(slot) of the batch that used it has been waited for by the caller, or
        # at least `depth` submits ago its read-back was queued before this forward's on out_stream)
        if self.count >= self.depth:
            compute.wait_event(self.out_ready[slot])
        logits = self.engine(self.dev_in[slot][:n], out=self.dev_out[slot][:n])
        self.in_free[slot].record(compute)
        self.fwd_done[slot].record(compute)
        with torch.cuda.stream(self.out_stream):
            self.out_stream.wait_event(self.fwd_done[slot])
            self.host_out[slot][:n].copy_(logits, non_blocking=True)
            self.out_ready[slot].record(self.out_stream)
        self.count += 1
        self.unread[slot] = True
        return slot

    def result(self, slot):
        """Wait for the logits of the batch submitted into ``slot``; the returned pinned tensor is
        valid until ``depth`` more batches have been submitted."""
        self.out_ready[slot].synchronize()
        self.unread[slot] = False
        return self.host_out[slot]
