"""CPU oracle for the R(2+1)D hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import this module, and only as the checker or
the timed CPU baseline.  Nothing under ``rnb_b200/`` imports it.

PARITY UNPINNED.  The arithmetic of this path lives in the third-party git
submodule ``irhum/R2Plus1D-PyTorch`` pinned at commit
76c674df225a54b7dae67e4f8509ed9df297b60c (reference ``.gitmodules:1-3``,
``.SUBMODULES.json:9-16``).  That submodule is absent from ``/root/reference``
and the reference ships no tests, golden vectors or checkpoints, so this file
is a *restatement* of the published architecture on top of ``torch.nn`` (fp32):

pinned by reference code
  * attribute names ``res2plus1d.conv1..conv5``, ``pool``, ``linear``
    (``models/r2p1d/model.py:55-62``, ``models/r2p1d/network.py:22-36,53-56``)
  * the ctor call shapes of every stage (``models/r2p1d/network.py:22-34``):
    stem ``SpatioTemporalConv(3, 64, [3,7,7], stride=[1,2,2], padding=[1,3,3])``
    and ``SpatioTemporalResLayer(cin, cout, 3, n, block_type=, downsample=)``
  * per-stage activation shapes (``models/r2p1d/model.py:29-33``)
  * ``AdaptiveAvgPool3d(1)``, ``view(-1, 512)``, ``Linear(512, num_classes)``
    (``models/r2p1d/network.py:35,41,56,60``)
  * layer-range semantics of ``R2Plus1DLayerNet`` / ``R2Plus1DLayerWrapper``
    (``models/r2p1d/network.py:9-60``), including the positional indexing of
    ``layer_sizes`` by *position in the range* (``network.py:20-34``)
  * preprocessing ``frames.float().permute(0, 2, 1, 3, 4)``
    (``models/r2p1d/model.py:149-150``)
  * video result ``logits.sum(axis=0).argmax()`` (``models/r2p1d/model.py:254,259``)
  * segment row ranges (``runner.py:149-151``)

assumed from the public submodule (not verifiable offline)
  * intermediate width ``floor(kt*kh*kw*cin*cout / (kh*kw*cin + kt*cout))``
  * ``bias=True`` on both factored convolutions
  * ``BatchNorm3d`` + ``ReLU`` between the spatial and the temporal convolution,
    nothing after the temporal one inside ``SpatioTemporalConv``
  * no BN/ReLU after the stem
  * residual block wiring ``bn1/relu1/bn2/outrelu`` and a shortcut made of
    ``SpatioTemporalConv(cin, cout, 1, stride=2)`` + ``BatchNorm3d``
  * torch's default initialisers

One deliberate deviation from the reference *as written*: the reference never
calls ``.eval()`` (only ``torch.no_grad()`` at ``runner.py:45``), so its
BatchNorm layers use batch statistics.  The contract of this build is
inference-mode BatchNorm, so every model made here is put in ``eval()``.
"""
import math

import torch
import torch.nn as nn

NUM_CLASSES = 400
FEATURES = 512
CLIP_SHAPE = (3, 8, 112, 112)

# input shape per start layer, leading clip count removed (model.py:29-33)
LAYER_INPUT_SHAPE = {
    1: (3, 8, 112, 112),
    2: (64, 8, 56, 56),
    3: (64, 8, 56, 56),
    4: (128, 4, 28, 28),
    5: (256, 2, 14, 14),
}
# output shape per end layer
LAYER_OUTPUT_SHAPE = {
    1: (64, 8, 56, 56),
    2: (64, 8, 56, 56),
    3: (128, 4, 28, 28),
    4: (256, 2, 14, 14),
    5: (NUM_CLASSES,),
}


def _triple(v):
    if isinstance(v, (list, tuple)):
        assert len(v) == 3
        return tuple(int(e) for e in v)
    return (int(v),) * 3


def intermediate_channels(cin, cout, kernel_size):
    """Width of the tensor between the spatial and the temporal convolution."""
    kt, kh, kw = _triple(kernel_size)
    return int(math.floor((kt * kh * kw * cin * cout) / (kh * kw * cin + kt * cout)))


class SpatioTemporalConv(nn.Module):
    """(2+1)D factorised convolution: 1 x kh x kw conv, BN, ReLU, kt x 1 x 1 conv."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, bias=True):
        super().__init__()
        kt, kh, kw = _triple(kernel_size)
        st, sh, sw = _triple(stride)
        pt, ph, pw = _triple(padding)
        mid = intermediate_channels(in_channels, out_channels, (kt, kh, kw))
        self.spatial_conv = nn.Conv3d(in_channels, mid, (1, kh, kw), stride=(1, sh, sw),
                                      padding=(0, ph, pw), bias=bias)
        self.bn = nn.BatchNorm3d(mid)
        self.relu = nn.ReLU()
        self.temporal_conv = nn.Conv3d(mid, out_channels, (kt, 1, 1), stride=(st, 1, 1),
                                       padding=(pt, 0, 0), bias=bias)

    def forward(self, x):
        return self.temporal_conv(self.relu(self.bn(self.spatial_conv(x))))


class SpatioTemporalResBlock(nn.Module):
    """Two factorised convolutions with an identity or strided-projection shortcut."""

    def __init__(self, in_channels, out_channels, kernel_size, downsample=False):
        super().__init__()
        self.downsample = downsample
        padding = kernel_size // 2
        if downsample:
            self.downsampleconv = SpatioTemporalConv(in_channels, out_channels, 1, stride=2)
            self.downsamplebn = nn.BatchNorm3d(out_channels)
            self.conv1 = SpatioTemporalConv(in_channels, out_channels, kernel_size,
                                            padding=padding, stride=2)
        else:
            self.conv1 = SpatioTemporalConv(in_channels, out_channels, kernel_size, padding=padding)
        self.bn1 = nn.BatchNorm3d(out_channels)
        self.relu1 = nn.ReLU()
        self.conv2 = SpatioTemporalConv(out_channels, out_channels, kernel_size, padding=padding)
        self.bn2 = nn.BatchNorm3d(out_channels)
        self.outrelu = nn.ReLU()

    def forward(self, x):
        res = self.relu1(self.bn1(self.conv1(x)))
        res = self.bn2(self.conv2(res))
        if self.downsample:
            x = self.downsamplebn(self.downsampleconv(x))
        return self.outrelu(x + res)


class SpatioTemporalResLayer(nn.Module):
    """``layer_size`` residual blocks; only the first may downsample."""

    def __init__(self, in_channels, out_channels, kernel_size, layer_size,
                 block_type=SpatioTemporalResBlock, downsample=False):
        super().__init__()
        self.block1 = block_type(in_channels, out_channels, kernel_size, downsample)
        self.blocks = nn.ModuleList(
            [block_type(out_channels, out_channels, kernel_size) for _ in range(layer_size - 1)])

    def forward(self, x):
        x = self.block1(x)
        for block in self.blocks:
            x = block(x)
        return x


_STAGE_CHANNELS = {2: (64, 64, False), 3: (64, 128, True), 4: (128, 256, True), 5: (256, 512, True)}


class R2Plus1DLayerNet(nn.Module):
    """Layers ``start_idx..end_idx`` (1 = stem ... 5 = conv5 + pool); network.py:9-41."""

    def __init__(self, start_idx, end_idx, num_classes, layer_sizes,
                 block_type=SpatioTemporalResBlock):
        super().__init__()
        self.end_idx = end_idx
        self._order = []
        for pos, layer_index in enumerate(range(start_idx, end_idx + 1)):
            if layer_index == 1:
                self.conv1 = SpatioTemporalConv(3, 64, [3, 7, 7], stride=[1, 2, 2], padding=[1, 3, 3])
                self._order.append('conv1')
            else:
                cin, cout, down = _STAGE_CHANNELS[layer_index]
                layer = SpatioTemporalResLayer(cin, cout, 3, layer_sizes[pos],
                                               block_type=block_type, downsample=down)
                name = 'conv%d' % layer_index
                setattr(self, name, layer)
                self._order.append(name)
                if layer_index == 5:
                    self.pool = nn.AdaptiveAvgPool3d(1)
                    self._order.append('pool')

    def forward(self, x, taps=None):
        for name in self._order:
            x = getattr(self, name)(x)
            if taps is not None:
                taps[name] = x
        return x.view(-1, FEATURES) if self.end_idx == 5 else x


class R2Plus1DLayerWrapper(nn.Module):
    """R2Plus1DLayerNet plus the classifier when the range ends at layer 5; network.py:44-60."""

    def __init__(self, start_idx, end_idx, num_classes, layer_sizes,
                 block_type=SpatioTemporalResBlock):
        super().__init__()
        self.res2plus1d = R2Plus1DLayerNet(start_idx, end_idx, num_classes, layer_sizes, block_type)
        self.end_idx = end_idx
        if end_idx == 5:
            self.linear = nn.Linear(FEATURES, num_classes)

    def forward(self, x, taps=None):
        x = self.res2plus1d(x, taps)
        return self.linear(x) if self.end_idx == 5 else x


class R2Plus1DClassifier(R2Plus1DLayerWrapper):
    """Whole network; ``layer_sizes`` has 4 entries here (model.py:171,177)."""

    def __init__(self, num_classes=NUM_CLASSES, layer_sizes=(2, 2, 2, 2),
                 block_type=SpatioTemporalResBlock):
        super().__init__(1, 5, num_classes, [0] + list(layer_sizes), block_type)


def build(start_idx=1, end_idx=5, num_classes=NUM_CLASSES, layer_sizes=None, seed=0,
          harden=False):
    """Seeded model in eval mode.

    ``layer_sizes`` follows ``R2P1DRunner`` (model.py:46-47): one entry per layer of the
    range *including the stem position*, default 2 each.  ``harden=True`` additionally
    randomises the BatchNorm affine parameters and running statistics (default init leaves
    them at identity, which makes logits nearly input-independent and BN folding untested).
    """
    if layer_sizes is None:
        layer_sizes = [2 for _ in range(start_idx, end_idx + 1)]
    gen_state = torch.random.get_rng_state()
    torch.manual_seed(seed)
    model = R2Plus1DLayerWrapper(start_idx, end_idx, num_classes, list(layer_sizes))
    if harden:
        g = torch.Generator().manual_seed(seed + 1)
        for mod in model.modules():
            if isinstance(mod, nn.BatchNorm3d):
                n = mod.num_features
                mod.weight.data = 0.75 + 0.5 * torch.rand(n, generator=g)
                mod.bias.data = 0.2 * torch.randn(n, generator=g)
                mod.running_mean.data = 0.1 * torch.randn(n, generator=g)
                mod.running_var.data = 0.6 + 0.8 * torch.rand(n, generator=g)
            elif isinstance(mod, (nn.Conv3d, nn.Linear)):
                # keep activations O(1) through 20 factorised convolutions
                mod.weight.data *= 1.7
                mod.bias.data = 0.05 * torch.randn(mod.bias.shape, generator=g)
    torch.random.set_rng_state(gen_state)
    model.eval()
    for p in model.parameters():
        p.requires_grad_(False)
    return model


def preprocess(frames_u8):
    """uint8 (N, T, C, H, W) loader output -> fp32 (N, C, T, H, W); model.py:149-150."""
    return frames_u8.float().permute(0, 2, 1, 3, 4)


def synthetic_clips(batch, seed=1234):
    """Seeded stand-in for NVVL output: uint8 (batch, 8, 3, 112, 112), uniform 0..255."""
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 256, (batch, 8, 3, 112, 112), dtype=torch.uint8, generator=g)


@torch.no_grad()
def forward(model, x, taps=None):
    return model(x.contiguous(), taps)


@torch.no_grad()
def forward_u8(model, frames_u8, taps=None):
    return model(preprocess(frames_u8).contiguous(), taps)


def video_class(logits):
    """Sum of per-clip logits, then argmax; model.py:254,259."""
    return int(logits.numpy().sum(axis=0).argmax())


def segment_ranges(rows, num_segments):
    """Row range of each segment; runner.py:149-151."""
    q, r = divmod(rows, num_segments)
    return [(q * s + min(s, r), q * (s + 1) + min(s + 1, r)) for s in range(num_segments)]


def segmented_rows(rows, num_segments):
    """Rows of the shared tensor that carries one segment; control.py:66."""
    return int(math.ceil(rows / num_segments))


def flops_per_clip(layer_sizes=(2, 2, 2, 2), num_classes=NUM_CLASSES):
    """2 x MACs of every Conv3d and the Linear for one 3x8x112x112 clip (zero-pad taps counted)."""
    model = build(1, 5, num_classes, [0] + list(layer_sizes))
    total = [0]

    def hook(mod, inp, out):
        if isinstance(mod, nn.Conv3d):
            k = mod.kernel_size[0] * mod.kernel_size[1] * mod.kernel_size[2]
            total[0] += 2 * out.numel() * mod.in_channels * k
        elif isinstance(mod, nn.Linear):
            total[0] += 2 * out.numel() * mod.in_features

    hs = [m.register_forward_hook(hook) for m in model.modules()
          if isinstance(m, (nn.Conv3d, nn.Linear))]
    forward(model, torch.zeros(1, *CLIP_SHAPE))
    for h in hs:
        h.remove()
    return total[0]
