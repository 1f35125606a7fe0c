"""The hot-path part of the reference's quantization/clip_utils.py: apply_clip (clip_utils.py:328-346) and the op
lookup helpers it needs (:292-307).  Calibration datasets and HF model loading are out of scope."""
import torch
import torch.nn as nn

from . import _lib as L


def get_named_linears(module):
    return {name: m for name, m in module.named_modules() if isinstance(m, nn.Linear)}


def get_op_by_name(module, op_name):
    for name, m in module.named_modules():
        if name == op_name:
            return m
    raise ValueError(f"Cannot find op {op_name} in module {module}")


def get_op_name(module, op):
    for name, m in module.named_modules():
        if m is op:
            return name
    raise ValueError(f"Cannot find op {op} in module {module}")


def clip_weight_(w, min_val, max_val):
    """In-place w = clamp(w.reshape(co, n_g, G), min_val, max_val) through bdq_apply_clip."""
    L.require_cuda(w, min_val, max_val)
    assert w.is_contiguous()
    ng = max_val.numel()
    assert min_val.numel() == ng and w.numel() % ng == 0
    lo = min_val.to(w.dtype).contiguous()
    hi = max_val.to(w.dtype).contiguous()
    with torch.cuda.device(w.device):
        rc = L.lib().bdq_apply_clip(L.ptr(w), L.ptr(lo), L.ptr(hi), L.i64(w.numel()), int(w.numel() // ng),
                                    L.dtype_code(w), L.stream_ptr(w.device))
    L.check(rc, "bdq_apply_clip")
    return w


@torch.no_grad()
def apply_clip(module, clip_results):
    if not isinstance(clip_results, list):
        clip_list = clip_results["clip"]
    else:
        clip_list = clip_results
    if len(clip_list) == 0:
        return
    if len(clip_list[0]) == 3:
        for name, max_val, min_val in clip_list:
            layer = get_op_by_name(module, name)
            layer.cuda()
            max_val = max_val.to(layer.weight.device)
            min_val = min_val.to(layer.weight.device)
            if not layer.weight.data.is_contiguous():
                layer.weight.data = layer.weight.data.contiguous()
            clip_weight_(layer.weight.data, min_val, max_val)
    else:
        # the reference's other branch is `raise 1` (clip_utils.py:346-347) -> TypeError
        raise TypeError("exceptions must derive from BaseException")
