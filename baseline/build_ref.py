"""Ship the UNMODIFIED reference to the GPU box: `baseline/_ref/` is git-ignored but travels with
the gpurun snapshot.  Run in the build container only (needs /root/reference); `__graft_entry__.build()` calls it.

What lands in baseline/_ref/:
  quantization/{quantizer,qlinear,qmodule,triton_kernels,autoclip,clip_utils}.py   byte-for-byte copies
  csrc/quantization/{gemv_cuda.cu,gemv_cuda.h,gemm_cuda_gen.cu,gemm_cuda.h,dequantize.cuh}   byte-for-byte copies
  awq_inference_engine.so   the reference's own `gemv_forward_cuda` / `gemm_forward_cuda` (inference/kernels/csrc/
      quantization/gemv_cuda.cu:275-346, gemm_cuda_gen.cu) compiled for sm_100 with the reference's nvcc flags
      (inference/kernels/setup.py:15-28) behind `baseline/ref_binding.cpp` -- the reference's pybind.cpp minus the
      attention / layernorm / rotary kernels that are out of scope, so `import awq_inference_engine` in the
      reference's qmodule.py resolves to the reference's kernels.
Nothing under bitdistiller_b200/ or in bench.py's product arm imports any of it: it is the parity pin and the perf bar
(tests/test_reference_gpu.py, tools/bench_vs_reference.py, bench.py `secondary.*.vs_ref*`).
"""
import filecmp
import os
import shutil
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
OUT = os.path.join(HERE, "_ref")

PY_FILES = ["quantizer.py", "qlinear.py", "qmodule.py", "triton_kernels.py", "autoclip.py", "clip_utils.py"]
C_FILES = ["gemv_cuda.cu", "gemv_cuda.h", "gemm_cuda_gen.cu", "gemm_cuda.h", "dequantize.cuh"]


def _copy(src, dst):
    os.makedirs(os.path.dirname(dst), exist_ok=True)
    if not (os.path.exists(dst) and filecmp.cmp(src, dst, shallow=False)):
        shutil.copyfile(src, dst)


def build(verbose=False):
    if not os.path.isdir(REF):
        print("baseline: /root/reference not present; using the prebuilt baseline/_ref as is")
        return False
    for f in PY_FILES:
        _copy(os.path.join(REF, "quantization", f), os.path.join(OUT, "quantization", f))
    for f in C_FILES:
        _copy(os.path.join(REF, "inference/kernels/csrc/quantization", f), os.path.join(OUT, "csrc/quantization", f))
    so = os.path.join(OUT, "awq_inference_engine.so")
    srcs = [os.path.join(OUT, "csrc/quantization/gemv_cuda.cu"), os.path.join(OUT, "csrc/quantization/gemm_cuda_gen.cu"),
            os.path.join(HERE, "ref_binding.cpp")]
    if os.path.exists(so) and all(os.path.getmtime(so) >= os.path.getmtime(s) for s in srcs):
        return True
    import torch
    from torch.utils import cpp_extension as ce
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    inc = [f"-I{p}" for p in ce.include_paths("cuda")] + [f"-I{sysconfig.get_paths()['include']}",
                                                           f"-I{os.path.join(OUT, 'csrc')}",
                                                           f"-I{os.path.join(OUT, 'csrc/quantization')}"]
    defs = ["-DTORCH_EXTENSION_NAME=awq_inference_engine", "-DTORCH_API_INCLUDE_EXTENSION_H",
            f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"]
    # the reference's nvcc flags (inference/kernels/setup.py:15-28); arch = the box's GPU
    nvflags = ["-O3", "-std=c++17", "-DENABLE_BF16", "-U__CUDA_NO_HALF_OPERATORS__", "-U__CUDA_NO_HALF_CONVERSIONS__",
               "-U__CUDA_NO_BFLOAT16_OPERATORS__", "-U__CUDA_NO_BFLOAT16_CONVERSIONS__",
               "-U__CUDA_NO_BFLOAT162_OPERATORS__", "-U__CUDA_NO_BFLOAT162_CONVERSIONS__", "--expt-relaxed-constexpr",
               "--expt-extended-lambda", "--use_fast_math", "-gencode", "arch=compute_100,code=sm_100",
               "-Xcompiler", "-fPIC"]
    bdir = os.path.join(OUT, "build")
    os.makedirs(bdir, exist_ok=True)
    objs, procs = [], []
    for s in srcs:   # the three translation units compile in parallel (each spends minutes in the PyTorch headers)
        o = os.path.join(bdir, os.path.basename(s) + ".o")
        cmd = ["nvcc"] + nvflags + defs + inc + ["-c", s, "-o", o]
        if verbose:
            print(" ".join(cmd))
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.DEVNULL if not verbose else None,
                                            stderr=subprocess.DEVNULL if not verbose else None)))
        objs.append(o)
    for cmd, pr in procs:
        if pr.wait() != 0:
            raise subprocess.CalledProcessError(pr.returncode, cmd)
    cmd = ["nvcc", "-shared", "-o", so] + objs + [f"-L{tlib}", "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda",
                                                  "-ltorch", "-ltorch_python", f"-Xlinker=-rpath={tlib}"]
    subprocess.check_call(cmd)
    return True


if __name__ == "__main__":
    ok = build(verbose="-v" in sys.argv)
    print("baseline/_ref:", "built" if ok else "not rebuilt", sorted(os.listdir(OUT)) if os.path.isdir(OUT) else None)
