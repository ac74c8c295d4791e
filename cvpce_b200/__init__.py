"""cvpce_b200: B200-native (sm_100a) proposal -> product classification hot path of laitalaj/cvpce.

Layout: csrc/ (CUDA kernels + the C ABI of include/cvpce_b200.h), _lib.py (ctypes loader), ops.py (tensor-in /
tensor-out wrappers), detection.py / classification.py / production.py / datautils.py / metrics.py / planograms.py
(host-side mirror of the reference interface for this path), patch.py (drop-in patchers for an installed reference), pipeline.py (batched path used by bench.py),
dist.py (one process per GPU: image data-parallel + gallery-sharded search).
"""
from . import _lib  # noqa: F401

__version__ = "0.1.0"
