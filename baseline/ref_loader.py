"""Loader for the UNMODIFIED reference shipped to the GPU box under baseline/_ref/ (baseline/build_ref.py): the
reference's quantization/*.py (CUDA eager + its Triton kernel) and its gemv_cuda.cu / gemm_cuda_gen.cu compiled for
sm_100 as `awq_inference_engine`.  Tests, tools and bench.py's `secondary` comparison only -- nothing in
bitdistiller_b200/ imports this."""
import importlib
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")


def have_ref():
    return os.path.exists(os.path.join(REF, "quantization", "quantizer.py")) and \
        os.path.exists(os.path.join(REF, "awq_inference_engine.so"))


def load():
    """-> dict(quantizer, qlinear, qmodule, triton_kernels, engine) of reference modules; RuntimeError if not shipped."""
    if not have_ref():
        raise RuntimeError("baseline/_ref is missing: run `python baseline/build_ref.py` in the build container "
                           "(needs /root/reference)")
    import torch  # noqa: F401  (libtorch must be loaded before the extension)
    for p in (os.path.join(REF, "quantization"), REF):
        if p not in sys.path:
            sys.path.insert(0, p)
    engine = importlib.import_module("awq_inference_engine")
    assert engine.__file__.startswith(REF), engine.__file__
    mods = {"engine": engine,
            "quantizer": importlib.import_module("quantizer"),          # qlinear.py does `from quantizer import *`
            "qlinear": importlib.import_module("qlinear"),
            "qmodule": importlib.import_module("quantization.qmodule"),  # relative import of .triton_kernels
            "triton_kernels": importlib.import_module("quantization.triton_kernels")}
    for m in ("quantizer", "qlinear", "qmodule", "triton_kernels"):
        assert mods[m].__file__.startswith(REF), mods[m].__file__
    return mods
