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


def append_str_prefix(x, prefix):
    """clip_utils.py:257-265: prefix every string of a nested list / tuple structure."""
    if isinstance(x, str):
        return prefix + x
    elif isinstance(x, tuple):
        return tuple([append_str_prefix(y, prefix) for y in x])
    elif isinstance(x, list):
        return [append_str_prefix(y, prefix) for y in x]
    else:
        return x


def get_blocks(model):
    from .pre_quant import get_blocks as _gb
    return _gb(model)


def get_calib_dataset(datasets="pileval", tokenizer=None, n_samples=128, block_size=1024):
    """clip_utils.py:32-48 dispatches on hub dataset names ("pile", "gsm8k", "code", ...) that need network access or the
    authors' local paths: that loader is outside the hot path (out of scope).  Here `datasets` may be a local
    text / jsonl file (one sample per line, jsonl key "text"); samples are tokenised, concatenated and cut into
    [1, block_size] blocks exactly like the reference loaders (clip_utils.py:72-77)."""
    import json
    import os
    if not (isinstance(datasets, str) and os.path.exists(datasets)):
        raise NotImplementedError(
            f"calibration corpus {datasets!r}: hub loaders are out of scope here; pass a local text/jsonl path, or give "
            "run_clip the token blocks directly (samples=...)")
    samples, n_run = [], 0
    with open(datasets) as f:
        for line in f:
            line = line.strip()
            if datasets.endswith(".jsonl") and line:
                line = json.loads(line)["text"].strip()
            enc = tokenizer.encode(line)
            if len(enc) > 512 or len(enc) == 0:
                continue
            samples.append(torch.tensor([enc]))
            n_run += 1
            if n_run == n_samples:
                break
    cat = torch.cat(samples, dim=1)
    n_split = cat.shape[1] // block_size
    print(f" * Split into {n_split} blocks")
    return [cat[:, i * block_size:(i + 1) * block_size] for i in range(n_split)]


def clip_weight_(w, min_val, max_val):
    """In-place w = clamp(w.reshape(co, n_g, G), min_val, max_val) through bdq_apply_clip."""
    L.require_cuda(w, min_val, max_val)
    assert w.is_contiguous()
    ng = max_val.numel()
    assert min_val.numel() == ng and w.numel() % ng == 0
    lo = min_val.to(w.dtype).contiguous()
    hi = max_val.to(w.dtype).contiguous()
    with L.on(w.device):
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
