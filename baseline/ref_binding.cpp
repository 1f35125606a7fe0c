// Binding for the UNMODIFIED reference kernels compiled into baseline/_ref/awq_inference_engine.so (test / bench
// baseline only).  Same module and function names as the reference's inference/kernels/csrc/pybind.cpp:9-13, minus the
// attention / layernorm / rotary entries that are out of scope, so the reference's qmodule.py imports it as is.
#include <pybind11/pybind11.h>
#include <torch/extension.h>
#include "quantization/gemm_cuda.h"
#include "quantization/gemv_cuda.h"

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
    m.def("gemm_forward_cuda", &gemm_forward_cuda, "reference 4-bit GEMM (gemm_cuda_gen.cu)");
    m.def("gemv_forward_cuda", &gemv_forward_cuda, "reference GEMV (gemv_cuda.cu)");
}
