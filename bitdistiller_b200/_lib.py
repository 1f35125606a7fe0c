"""ctypes binding of libbdq.so (the C ABI in include/bdq.h).

There is NO fallback: if the shared library is missing or a tensor is not on a CUDA device the call raises.
PyTorch is used only for device memory and streams.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libbdq.so")

F32, BF16, F16 = 0, 1, 2
SEM_CPU, SEM_CUDA = 0, 1
ROUND_HALF_EVEN, ROUND_HALF_AWAY = 0, 1
QUANT_INT, QUANT_NF3 = 0, 1

_DTYPES = {torch.float32: F32, torch.bfloat16: BF16, torch.float16: F16}

_lib = None


class BdqError(RuntimeError):
    pass


def semantics():
    """Eager-device rounding semantics to reproduce: `tensor / python_scalar` is a true division in
    CPU eager and a multiply by the fp32 reciprocal in CUDA eager, so the reference's scales differ by 1 ulp in ~1/3 of
    the groups between the two devices.  Default = `cuda`: this package runs on CUDA tensors only, and every real flow of
    the reference (QAT, run_clip, real_quantize_model_weight) quantises on the GPU, so a user swapping the import gets the
    bits the reference produces on that same GPU (pinned on the B200 against the unmodified reference in CUDA eager,
    tests/test_reference_gpu.py).  BDQ_SEMANTICS=cpu reproduces the reference's CPU results (pinned by tests/golden/*.npz,
    which can only be generated on the GPU-less build container); tests/conftest.py selects it for the golden tests."""
    v = os.environ.get("BDQ_SEMANTICS", "cuda").lower()
    if v not in ("cuda", "cpu"):
        raise BdqError(f"BDQ_SEMANTICS must be 'cuda' or 'cpu', got {v!r}")
    return SEM_CUDA if v == "cuda" else SEM_CPU


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            raise BdqError(
                f"{_SO} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(or `make -C bitdistiller_b200/csrc`). bitdistiller_b200 has no CPU / PyTorch fallback.")
        _lib = ctypes.CDLL(_SO)
        _lib.bdq_last_error.restype = ctypes.c_char_p
        _lib.bdq_launch_count.restype = ctypes.c_int64
    return _lib


def launch_count():
    return int(lib().bdq_launch_count())


def check(rc, what):
    if rc != 0:
        msg = lib().bdq_last_error().decode("utf-8", "replace")
        raise BdqError(f"{what} failed (code {rc}): {msg}")


def dtype_code(t):
    try:
        return _DTYPES[t.dtype]
    except KeyError:
        raise TypeError(f"bitdistiller_b200 supports float32 / bfloat16 / float16 tensors, got {t.dtype}") from None


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise BdqError("bitdistiller_b200 runs on CUDA tensors only (no CPU fallback); got a tensor on "
                           f"{t.device}. Move it to the GPU first.")


def ptr(t):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


class _Noop:
    def __enter__(self):
        return None

    def __exit__(self, *exc):
        return False


_NOOP = _Noop()


def on(device):
    """`with L.on(t.device):` makes `device` current for a library call.  torch.cuda.device() costs several microseconds of
    Python per call (prefill issues hundreds of 20-50 us GEMMs per forward), so when the device already is the current one
    this is a no-op."""
    idx = device.index if isinstance(device, torch.device) else device
    if idx is None or idx == torch.cuda.current_device():
        return _NOOP
    return torch.cuda.device(device)


def stream_ptr(device=None):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def i64(v):
    return ctypes.c_int64(int(v))
