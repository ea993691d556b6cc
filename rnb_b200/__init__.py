"""rnb_b200: the R(2+1)D hot path of the RnB benchmark, rebuilt for NVIDIA B200 (sm_100a).

``rnb_b200.steps`` holds the pipeline step classes a ``config/*.json`` names;
``rnb_b200.engine.R2P1DEngine`` is the torch-facing model handle; ``rnb_b200.ops`` wraps the
stand-alone kernels.  All arithmetic lives in ``csrc/librnb_b200.so`` (C ABI in
``include/rnb_b200.h``); there is no fallback path.
"""
__version__ = '0.1.0'
