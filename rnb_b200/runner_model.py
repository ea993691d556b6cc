"""Step plugin interface of the RnB benchmark, as the reference defines it.

This is the drop-in boundary: RnB's ``runner.py`` instantiates a step with
``model_class(device, **model_kwargs)`` (reference ``runner.py:47-48``), calls the static
``output_shape()`` in the main process to size shared tensors (``control.py:125-127``) and then
calls the instance once per request (``runner.py:125-126``).  The class below restates the
contract of reference ``runner_model.py:1-81``; when the reference tree is importable the step
classes in :mod:`rnb_b200.steps` derive from the reference's own ``RunnerModel`` instead, so that
``isinstance`` checks in a patched harness keep working.
"""

try:  # inside an RnB checkout the reference's own base class wins
    from runner_model import RunnerModel as _ReferenceRunnerModel  # type: ignore
    if _ReferenceRunnerModel.__module__ == __name__:  # guard against importing ourselves
        raise ImportError
except Exception:  # noqa: BLE001 - any failure means "not inside an RnB checkout"
    _ReferenceRunnerModel = None


class _RunnerModel:
    """A pipeline step: ``__init__(device, **kwargs)``, shapes, and ``__call__``."""

    def __init__(self, device):
        pass

    def input_shape(self):
        """Nested tuple of per-tensor shapes this step expects, or None (no tensor input)."""
        raise NotImplementedError

    @staticmethod
    def output_shape():
        """Nested tuple of per-tensor shapes this step emits, or None (no tensor output).

        Called without an instance, before any process is spawned, so it must not touch CUDA.
        """
        raise NotImplementedError

    def __call__(self, tensors, non_tensors, time_card):
        """Process one request.

        ``tensors`` is a tuple of device tensors (or None), ``non_tensors`` any object (or None),
        ``time_card`` the request's timing record.  Returns the same triple for the next step;
        returning ``None`` as the time card means "nothing to emit yet".
        """
        raise NotImplementedError


RunnerModel = _ReferenceRunnerModel if _ReferenceRunnerModel is not None else _RunnerModel
