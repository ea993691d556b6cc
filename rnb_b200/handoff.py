"""Tensor hand-off between pipeline steps: segment arithmetic and fenced (peer) copies.

Restates the data-moving lines of the reference's worker loop (``runner.py:63-72,90-121,138-173``)
and of ``control.py:49-69`` on top of ``rnb_peer_copy``.  The reference publishes a shared tensor
right after *enqueueing* the copy and releases it the same way (no stream sync on either side,
SURVEY.md section 5); the helpers here return only after the copy has completed on its stream, so
publishing or releasing afterwards is safe across processes.
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


def publish(shared_tensor, segment):
    """Producer side: copy ``segment`` into the leading rows of ``shared_tensor`` and wait.

    Returns the number of valid rows (what the reference stores in ``valid_batch_sizes``,
    runner.py:170-171).  ``shared_tensor`` may live on another GPU.
    """
    n = int(segment.shape[0])
    if n:
        ops.peer_copy(shared_tensor[:n], segment.contiguous())
        torch.cuda.current_stream(shared_tensor.device).synchronize()
    return n


def fetch(placeholder, shared_tensor, valid_rows):
    """Consumer side: pull the valid rows into the runner-owned placeholder and wait.

    After this returns the shared tensor may be released (``event.set()``, runner.py:117).
    """
    n = int(valid_rows)
    if n:
        ops.peer_copy(placeholder[:n], shared_tensor[:n])
        torch.cuda.current_stream(placeholder.device).synchronize()
    return placeholder[:n]
