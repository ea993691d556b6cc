"""Tensor hand-off between pipeline steps: shared buffers, segment arithmetic, fenced (peer) copies.

Restates the data-moving lines of the reference's worker loop (``runner.py:63-72,90-121,138-173``)
and the buffer type behind them (``control.py:19-69``) on top of ``rnb_peer_copy_fenced``.

What differs from the reference, on purpose:

* **Ordering.**  The reference publishes a shared tensor right after *enqueueing* the copy into it
  and releases it right after enqueueing the copy out of it, with no stream synchronisation on
  either side although producer and consumer are different processes.  Here
  every buffer carries two CUDA events that cross the process boundary (``cudaEventInterprocess``):
  ``ready`` is recorded by the producer behind its copy and waited for -- on the GPU -- by the
  consumer's copy; ``free`` is recorded by the consumer behind its copy and waited for by the
  producer's next copy into the buffer.  The ``multiprocessing.Event`` of the reference stays as
  the host-side flow control; no host thread waits for a GPU.
* **Wire format.**  ``control.py:41`` fixes the shared tensors to fp32.  ``TensorEvent`` takes a dtype
  per tensor, so a loader can ship the uint8 frames NVVL produced (301 056 B per clip instead of
  1 204 224 B) and the network's first kernel converts them (``RNB_FMT_U8_NTCHW``).
"""
import math

import torch

from . import ops


def segment_ranges(rows, num_segments):
    """Row range of every segment: remainders go to the first segments (runner.py:149-151)."""
    q, r = divmod(rows, num_segments)
    return [(q * s + min(s, r), q * (s + 1) + min(s + 1, r)) for s in range(num_segments)]


def get_segmented_shapes(shapes, num_segments):
    """Shapes of the shared tensors that carry one segment each (control.py:49-69)."""
    if shapes is None or num_segments == 1:
        return shapes
    new_shapes = []
    for shape in shapes:
        batch_size = shape[0]
        if num_segments > batch_size:
            raise ValueError('num_segments %d must be <= tensor batch size %d'
                             % (num_segments, batch_size))
        new_shapes.append((int(math.ceil(batch_size / num_segments)),) + tuple(shape[1:]))
    return new_shapes


class TensorEvent:
    """Shared tensors + release flag + valid row counts: ``control.TensorEvent`` with a dtype.

    Same attributes as the reference class (``tensors``, ``event``, ``valid_batch_sizes``;
    control.py:40-46), so the reference's worker loop can use it unchanged.  ``dtype`` is one
    ``torch.dtype`` for all tensors or one per tensor.  On a CUDA device the object also carries the
    two interprocess CUDA events described in the module docstring; they travel through
    ``torch.multiprocessing`` pickling like the tensors do.
    """

    def __init__(self, shapes, device, dtype=torch.float32, ctx=None):
        # `ctx`: a multiprocessing context for the host-side primitives (default: the process-wide
        # start method, which RnB sets to 'spawn' before it builds any of these, benchmark.py:132)
        if ctx is None:
            import torch.multiprocessing as ctx
        Array, Event = ctx.Array, ctx.Event
        device = torch.device(device)
        dtypes = list(dtype) if isinstance(dtype, (list, tuple)) else [dtype] * len(shapes)
        if len(dtypes) != len(shapes):
            raise ValueError('one dtype per shape expected')
        self.tensors = tuple(torch.empty(*shape, dtype=dt, device=device)
                             for shape, dt in zip(shapes, dtypes))
        self.event = Event()
        self.event.set()
        self.valid_batch_sizes = Array('i', len(shapes))
        self.ready = self.free = None
        if device.type == 'cuda':
            with torch.cuda.device(device):
                self.ready = torch.cuda.Event(enable_timing=False, interprocess=True)
                self.free = torch.cuda.Event(enable_timing=False, interprocess=True)
                # torch creates the CUDA event lazily, on whatever device is current at its first
                # use -- which would be the pickling in the main process, on device 0.  Record both
                # once here, on THIS device's (idle) stream: the events then belong to the buffer's
                # device, and a first wait on either returns at once.
                s = torch.cuda.current_stream(device)
                self.ready.record(s)
                self.free.record(s)


_copy_streams = {}


def _buffer_stream(buffer_device, work_stream):
    """The stream a copy into / out of a shared buffer runs on.

    CUDA events can only be recorded on a stream of the device they were created on, and a
    buffer's ``ready`` / ``free`` events live on the buffer's device.  A step on the same GPU uses
    its own stream; a step on another GPU uses a private stream on the buffer's device (one per
    process and device), which a peer copy can run on just as well.
    """
    if work_stream.device == buffer_device:
        return work_stream
    key = buffer_device.index
    if key not in _copy_streams:
        _copy_streams[key] = torch.cuda.Stream(device=buffer_device)
    return _copy_streams[key]


def publish(shared, segment, index=0):
    """Producer side (runner.py:163-173): copy ``segment`` into the leading rows of a shared buffer.

    ``shared`` is a ``TensorEvent`` (this module's or the reference's) or a bare tensor.  The copy
    is ordered behind the kernels that produced the segment (the current stream of its device) and
    -- when the buffer carries events -- behind the previous consumer's copy out of the buffer.
    With events the call returns without waiting for the GPU; without them (reference
    ``TensorEvent``, bare tensor) it waits for the copy, so that enqueueing the ``Signal``
    afterwards is safe either way.  Returns the number of valid rows (``valid_batch_sizes``,
    runner.py:170-171).
    """
    tensor = shared.tensors[index] if hasattr(shared, 'tensors') else shared
    n = int(segment.shape[0])
    work = torch.cuda.current_stream(segment.device)
    stream = _buffer_stream(tensor.device, work)
    free = getattr(shared, 'free', None)
    ready = getattr(shared, 'ready', None)
    if free is not None:
        stream.wait_event(free)
    if n:
        seg = segment if segment.is_contiguous() else segment.contiguous()
        ops.peer_copy(tensor[:n], seg, src_stream=work, on=stream)
    elif stream is not work:
        stream.wait_stream(work)
    if hasattr(shared, 'valid_batch_sizes'):
        shared.valid_batch_sizes[index] = n
    if ready is not None:
        ready.record(stream)
    else:
        stream.synchronize()
    return n


def fetch(placeholder, shared, valid_rows=None, index=0):
    """Consumer side (runner.py:107-117): pull the valid rows into the runner-owned placeholder.

    Ordered behind the producer's copy (``ready``) and behind whatever the consumer's current
    stream still does with the placeholder; the consumer's stream continues only after the rows
    have arrived.  With events the buffer may be released (``event.set()``, runner.py:117) as soon
    as this returns: the producer's next copy waits for ``free`` on the GPU.  Without events the
    copy is ordered behind the source device's current stream and the call waits for it.
    """
    tensor = shared.tensors[index] if hasattr(shared, 'tensors') else shared
    if valid_rows is None:
        valid_rows = shared.valid_batch_sizes[index]
    n = int(valid_rows)
    work = torch.cuda.current_stream(placeholder.device)
    stream = _buffer_stream(tensor.device, work)
    ready = getattr(shared, 'ready', None)
    free = getattr(shared, 'free', None)
    if ready is not None:
        stream.wait_event(ready)
    if stream is not work:
        stream.wait_stream(work)  # the placeholder's previous readers come first
    if n:
        # with events `ready` is the fence behind the producer (its stream belongs to another
        # process); without them the producer is taken to be the source device's current stream
        ops.peer_copy(placeholder[:n], tensor[:n], src_stream=stream if ready is not None else None,
                      on=stream)
    if free is not None:
        free.record(stream)
    if stream is not work:
        work.wait_stream(stream)
    if ready is None:
        work.synchronize()
    return placeholder[:n]
