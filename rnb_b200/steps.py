"""RnB pipeline steps for R(2+1)D, backed by the sm_100a kernels in ``librnb_b200.so``.

Drop-in counterparts of the classes in the reference's ``models/r2p1d/model.py`` and
``batcher.py``: same class names, constructor keywords, shapes and return conventions, selected
from a ``config/*.json`` pipeline file by dotted path (e.g. ``rnb_b200.steps.R2P1DRunner``).
Importing this module touches neither CUDA nor the shared library, because RnB's main process
imports step classes just to call ``output_shape()`` (reference control.py:125-127).

Differences from the reference, all deliberate:
  * inference-mode BatchNorm (the reference never calls ``.eval()``);
  * the NVVL video decoder is not available, so the loader steps serve seeded synthetic uint8
    clips of the same shape and clip-count distribution (models/r2p1d/sampler.py:58-59);
  * weights come from ``checkpoint=`` (a ``torch.save`` file with a ``state_dict`` entry, as at
    models/r2p1d/model.py:50-63), ``state_dict=`` or, by default, a seeded random init;
  * ``Batcher`` copies into its own buffer (the reference's stashed slices alias the runner's
    placeholder, batcher.py:21-28 + runner.py:112-114).
"""
import os
import random
import sys
from itertools import count

import numpy as np
import torch

from .runner_model import RunnerModel
from .timecard import TimeCard, TimeCardList

try:
    from selector import QueueSelector  # type: ignore  # reference selector.py:1-7
except Exception:  # noqa: BLE001
    class QueueSelector:
        def __init__(self, num_queues):
            pass

        def select(self, tensors, non_tensors, time_card):
            raise NotImplementedError

try:
    from video_path_provider import VideoPathIterator  # type: ignore  # video_path_provider.py:1-14
except Exception:  # noqa: BLE001
    class VideoPathIterator:
        def __init__(self):
            pass

        def __iter__(self):
            raise NotImplementedError

# Rows of the tensors handed between steps.  The reference hard-codes 15 clips for loader
# outputs (model.py:158, batcher.py:15) and 10 for runner outputs (model.py:80); one cap is used
# here so that a 15-clip video fits everywhere.  Batch sweeps raise it via the environment,
# because output_shape() is static and evaluated in the main process.
MAX_CLIPS = int(os.environ.get('RNB_B200_MAX_CLIPS', '15'))
NUM_CLASSES = 400
# What travels between the loader and the runner.  'fp32' is the reference's wire: fp32
# (n, 3, 8, 112, 112), 1.2 MB per clip (control.py:41, models/r2p1d/model.py:149-150).  'uint8' ships
# the frames as the decoder produced them, uint8 (n, 8, 3, 112, 112), a quarter of the bytes, and
# the network's first kernel does the float()/permute; it needs shared buffers that know their
# dtype (rnb_b200.handoff.TensorEvent, installed by scripts/run_rnb.py).  Static for the same reason
# MAX_CLIPS is: output_shape() / output_dtype() are evaluated in RnB's main process.
WIRE = os.environ.get('RNB_B200_WIRE', 'fp32')
if WIRE not in ('fp32', 'uint8'):
    raise ValueError("RNB_B200_WIRE must be 'fp32' or 'uint8', got %r" % WIRE)


def _clip_tensor_shape():
    return (MAX_CLIPS, 8, 3, 112, 112) if WIRE == 'uint8' else (MAX_CLIPS, 3, 8, 112, 112)


def _clip_tensor_dtype():
    return torch.uint8 if WIRE == 'uint8' else torch.float32


def _default_prepare_batches(max_batch):
    """Clip counts RnB's pipelines produce: 1 or 15 clips per video (sampler.py:58-59), 3-5 after
    3-way segmenting (control.py:66), 3 from Batcher(batch=3) (config/rnb.json:15), the nominal 10
    of the shape tables (model.py:29-33) and the configured cap."""
    return sorted(b for b in {1, 3, 4, 5, 10, 15, MAX_CLIPS} if b <= max_batch)


def _resolve_state_dict(start_index, end_index, num_classes, layer_sizes, checkpoint, state_dict,
                        seed):
    from .engine import random_state_dict
    if state_dict is not None:
        return state_dict
    if checkpoint is not None:
        ckpt = torch.load(checkpoint, map_location='cpu')
        return ckpt['state_dict'] if 'state_dict' in ckpt else ckpt
    return random_state_dict(start_index, end_index, num_classes, layer_sizes, seed)


class _SyntheticClips:
    """Seeded stand-in for ``nvvl.RnBLoader``: uint8 (n, 8, 3, 112, 112) frames on the GPU."""

    def __init__(self, device, num_clips_population, num_clips_weights, seed):
        self.population = list(num_clips_population)
        self.weights = list(num_clips_weights)
        self.rng = random.Random(seed)
        g = torch.Generator(device='cpu').manual_seed(seed)
        rows = max(self.population)
        self.pool = torch.randint(0, 256, (rows, 8, 3, 112, 112), dtype=torch.uint8,
                                  generator=g).to(device)

    def next(self):
        n = self.rng.choices(self.population, self.weights)[0]
        return self.pool[:n]


class R2P1DRunner(RunnerModel):
    """Any connected range of R(2+1)D layers (1-indexed ``start_index..end_index``)."""
    # expected input per start layer, with the reference's nominal 10 clips (model.py:29-33)
    input_dict = {1: (10, 3, 8, 112, 112),
                  2: (10, 64, 8, 56, 56),
                  3: (10, 64, 8, 56, 56),
                  4: (10, 128, 4, 28, 28),
                  5: (10, 256, 2, 14, 14)}

    def __init__(self, device, start_index=1, end_index=5, num_classes=NUM_CLASSES,
                 layer_sizes=None, block_type=None, dtype='bf16', checkpoint=None, state_dict=None,
                 seed=0, max_batch=64, warmup=3, prepare_batches=None):
        super(R2P1DRunner, self).__init__(device)
        from .engine import R2P1DEngine
        if start_index < 1:
            print('[ERROR] Wrong layer index for the starting layer! The start_index (%d) should '
                  'be more than or equal to 1.' % start_index)
            sys.exit()
        elif end_index > 5:
            print('[ERROR] Wrong layer index for the ending layer! The end_index (%d) should be '
                  'less than or equal to 5.' % end_index)
            sys.exit()
        if layer_sizes is None:
            layer_sizes = [2 for _ in range(start_index, end_index + 1)]
        self.start_index = start_index
        self.end_index = end_index
        sd = _resolve_state_dict(start_index, end_index, num_classes, layer_sizes, checkpoint,
                                 state_dict, seed)
        self.model = R2P1DEngine(device, sd, start_index, end_index, num_classes, layer_sizes,
                                 dtype=dtype, max_batch=max(max_batch, MAX_CLIPS))
        # Kernel selection for every clip count the pipeline can hand us, here rather than inside
        # the first live request of each size (the reference pays for cudnn.benchmark there,
        # runner.py:25), then a few warm-up inferences as the reference does (model.py:66-71).
        if prepare_batches is None:
            prepare_batches = _default_prepare_batches(self.model.max_batch)
        self.model.prepare(*prepare_batches)
        stream = torch.cuda.current_stream()
        tmp = torch.randn(*self.input_shape()[0], dtype=torch.float32, device=device)
        for _ in range(warmup):
            _ = self.model(tmp)
            stream.synchronize()

    def input_shape(self):
        return (self.input_dict[self.start_index],)

    @staticmethod
    def output_shape():
        # like the reference (model.py:77-80) this describes a range that ends at layer 5
        return ((MAX_CLIPS, NUM_CLASSES),)

    def __call__(self, tensors, non_tensors, time_card):
        tensor = tensors[0]
        return (self.model(tensor),), None, time_card


class R2P1DVideoPathIterator(VideoPathIterator):
    """Endless stream of synthetic video names (the reference walks a Kinetics-400 directory)."""

    def __init__(self):
        super(R2P1DVideoPathIterator, self).__init__()
        self.videos_iter = ('synthetic/video_%07d.mp4' % i for i in count())

    def __iter__(self):
        return self.videos_iter


class R2P1DLoader(RunnerModel):
    """Produces the clip tensor of one video: fp32 (n, 3, 8, 112, 112) like the reference, or the
    raw uint8 (n, 8, 3, 112, 112) frames when ``RNB_B200_WIRE=uint8``."""

    def __init__(self, device, num_clips_population=[1, 15], num_clips_weights=[10, 1], seed=0):
        super(R2P1DLoader, self).__init__(device)
        self.device = torch.device(device)
        self.clips = _SyntheticClips(self.device, num_clips_population, num_clips_weights, seed)
        self.out = torch.empty((max(MAX_CLIPS, max(num_clips_population)), 3, 8, 112, 112),
                               dtype=torch.float32, device=self.device)
        from . import ops
        self._preprocess = ops.preprocess_u8
        # warm up (the reference decodes three sample videos here, model.py:127-138)
        for _ in range(3):
            self._preprocess(self.clips.pool[:1], out=self.out[:1])
        torch.cuda.current_stream().synchronize()

    def __call__(self, tensors, non_tensors, time_card):
        frames = self.clips.next()
        time_card.num_clips = frames.shape[0]
        if WIRE == 'uint8':
            return (frames,), None, time_card
        # float() + permute(0, 2, 1, 3, 4) of model.py:149-150, written contiguously in one pass
        frames = self._preprocess(frames, out=self.out[:frames.shape[0]])
        return (frames,), None, time_card

    def input_shape(self):
        return None

    @staticmethod
    def output_shape():
        return (_clip_tensor_shape(),)

    @staticmethod
    def output_dtype():
        return (_clip_tensor_dtype(),)


class R2P1DSingleStep(RunnerModel):
    """Loader and whole network in one step: no tensor hand-off between processes."""
    input_tensor_shape = (10, 3, 8, 112, 112)

    def __init__(self, device, num_classes=NUM_CLASSES, layer_sizes=[2, 2, 2, 2], block_type=None,
                 num_clips_population=[1, 15], num_clips_weights=[10, 1], dtype='bf16',
                 checkpoint=None, state_dict=None, seed=0, max_batch=64, warmup=3):
        super(R2P1DSingleStep, self).__init__(device)
        from .engine import R2P1DEngine
        self.device = torch.device(device)
        full_sizes = [2] + list(layer_sizes)  # the classifier takes 4 sizes, the runner form 5
        sd = _resolve_state_dict(1, 5, num_classes, full_sizes, checkpoint, state_dict, seed)
        self.model = R2P1DEngine(device, sd, 1, 5, num_classes, full_sizes, dtype=dtype,
                                 max_batch=max(max_batch, MAX_CLIPS))
        self.clips = _SyntheticClips(self.device, num_clips_population, num_clips_weights, seed)
        self.model.prepare(*sorted(set(num_clips_population)))
        stream = torch.cuda.current_stream()
        for _ in range(warmup):
            _ = self.model(self.clips.pool)
            stream.synchronize()

    def __call__(self, tensors, non_tensors, time_card):
        frames = self.clips.next()
        time_card.num_clips = frames.shape[0]
        # the uint8 -> float conversion and the permute are fused into the network's first kernel
        return (self.model(frames),), None, time_card

    def input_shape(self):
        return None

    @staticmethod
    def output_shape():
        return ((MAX_CLIPS, NUM_CLASSES),)


class R2P1DVideoClassifier(RunnerModel):
    """Loader, whole network and aggregation in one final step; emits the video's class id.

    The per-clip logits never leave the GPU: the classifier kernel also sums over the clips and
    takes the arg-max (what R2P1DAggregator does on the host).  The class id lands in a pinned host
    tensor that is valid once the runner has synchronised its stream (runner.py:127-128); it is
    returned as the non-tensor output.
    """

    def __init__(self, device, num_classes=NUM_CLASSES, layer_sizes=[2, 2, 2, 2], block_type=None,
                 num_clips_population=[1, 15], num_clips_weights=[10, 1], dtype='bf16',
                 checkpoint=None, state_dict=None, seed=0, max_batch=64, warmup=3):
        super(R2P1DVideoClassifier, self).__init__(device)
        from .engine import R2P1DEngine
        self.device = torch.device(device)
        full_sizes = [2] + list(layer_sizes)
        sd = _resolve_state_dict(1, 5, num_classes, full_sizes, checkpoint, state_dict, seed)
        self.model = R2P1DEngine(device, sd, 1, 5, num_classes, full_sizes, dtype=dtype,
                                 max_batch=max(max_batch, MAX_CLIPS))
        self.clips = _SyntheticClips(self.device, num_clips_population, num_clips_weights, seed)
        self.cls_host = torch.empty((), dtype=torch.int32).pin_memory()
        self.model.prepare(*sorted(set(num_clips_population)))
        stream = torch.cuda.current_stream()
        for _ in range(warmup):
            self.model.classify(self.clips.pool)
            stream.synchronize()

    def __call__(self, tensors, non_tensors, time_card):
        frames = self.clips.next()
        time_card.num_clips = frames.shape[0]
        _, _, cls = self.model.classify(frames)
        self.cls_host.copy_(cls, non_blocking=True)
        return None, self.cls_host, time_card

    def input_shape(self):
        return None

    @staticmethod
    def output_shape():
        return None


class R2P1DAggregator(RunnerModel):
    """Sums per-clip logits of a video (over its segments) and picks the class."""

    def __init__(self, device, aggregate=1):
        super(R2P1DAggregator, self).__init__(device)
        # number of segments per video; equals num_segments of the producing step
        self.aggregate = aggregate
        self.results = {}

    @staticmethod
    def _clip_sum(tensor):
        if tensor.is_cuda:
            from . import ops
            total, _ = ops.logits_sum_argmax(tensor)
            return total.cpu().numpy()
        # the reference config places this step on the CPU (gpus: [-1]); host tensors stay there
        return tensor.numpy().sum(axis=0)

    def __call__(self, tensors, non_tensors, time_card):
        result = self._clip_sum(tensors[0])
        if self.aggregate == 1:
            return None, result.argmax(), time_card
        if time_card.id not in self.results:
            self.results[time_card.id] = (result, [time_card])
            return None, None, None
        prev_result, prev_time_cards = self.results[time_card.id]
        result = prev_result + result
        time_cards = prev_time_cards + [time_card]
        if len(time_cards) < self.aggregate:
            self.results[time_card.id] = (result, time_cards)
            return None, None, None
        del self.results[time_card.id]
        return None, result.argmax(), TimeCard.merge(time_cards)

    def input_shape(self):
        return ((MAX_CLIPS, NUM_CLASSES),)

    @staticmethod
    def output_shape():
        return None


class LargeSmallSelector(QueueSelector):
    """15-clip videos go to queue 1, everything else to queue 0 (model.py:288-296)."""

    def __init__(self, num_queues):
        assert num_queues == 2

    def select(self, tensors, non_tensors, time_card):
        if hasattr(time_card, 'num_clips') and time_card.num_clips == 15:
            return 1
        return 0


class Batcher(RunnerModel):
    """Collects ``batch`` requests and emits their clips as one contiguous tensor."""

    def __init__(self, device, batch=1):
        super(Batcher, self).__init__(device)
        self.device = torch.device(device)
        self.batch = batch
        self.rows = 0
        self.stacked_time_cards = []
        self.buffer = None
        if batch > 1:
            self.buffer = torch.empty(self.output_shape()[0], dtype=_clip_tensor_dtype(),
                                      device=self.device)

    @staticmethod
    def output_shape():
        return (_clip_tensor_shape(),)

    @staticmethod
    def output_dtype():
        return (_clip_tensor_dtype(),)

    def __call__(self, tensors, non_tensors, time_card):
        if self.batch <= 1:
            return tensors, non_tensors, time_card
        from . import ops
        tensor = tensors[0]
        n = tensor.shape[0]
        if self.rows + n > self.buffer.shape[0]:
            raise ValueError('Batcher: %d clips do not fit the %d-row batch buffer'
                             % (self.rows + n, self.buffer.shape[0]))
        # copy now: the runner reuses the memory behind `tensor` for the next request
        ops.batch_gather([tensor.contiguous()], self.buffer[self.rows:])
        self.rows += n
        self.stacked_time_cards.append(time_card)
        if len(self.stacked_time_cards) < self.batch:
            return None, None, None
        tensor_batch = self.buffer[:self.rows]
        time_card_list = TimeCardList(self.stacked_time_cards)
        self.rows = 0
        self.stacked_time_cards = []
        return (tensor_batch,), None, time_card_list
