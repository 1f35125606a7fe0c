"""bitdistiller_b200: B200-native (sm_100a) kernels behind BitDistiller's quantization API.

Module names mirror the reference's quantization/ directory so that swapping the import path is a drop-in:
quantizer, qlinear, qmodule, autoclip, clip_utils, triton_kernels, awq_inference_engine.
"""
from . import _lib  # noqa: F401

__all__ = ["quantizer", "qlinear", "qmodule", "autoclip", "clip_utils", "triton_kernels", "awq_inference_engine"]
