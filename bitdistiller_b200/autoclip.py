"""Drop-in for the hot-path part of the reference's quantization/autoclip.py: auto_2clip_layer (autoclip.py:13-83) and
auto_clip_block (:86-107).  One kernel launch (bdq_autoclip) evaluates every candidate of every (output channel, group)
pair; the reference loops 100 times over a [oc_batch, 512, ci] broadcast product.  run_clip / the CLI (HF model +
dataset loop) are out of scope."""
import gc

import torch
import torch.nn as nn

from . import _lib as L
from .clip_utils import apply_clip  # noqa: F401  (autoclip.py does `from clip_utils import *`)
from .quantizer import pseudo_quantize_n2f3_tensor, pseudo_quantize_tensor  # noqa: F401


@torch.no_grad()
def auto_2clip_layer(w, input_feat, n_bit, q_config, n_grid=20, max_shrink=0.5, n_sample_token=512, return_info=False):
    assert w.dim() == 2
    group_size = q_config["q_group_size"] if q_config["q_group_size"] > 0 else w.shape[1]
    quant_type = q_config["quant_type"]
    if quant_type not in ("int", "nf3"):
        raise ValueError(f"Has no support {quant_type}. Valid quant_type:[int, nf3]")
    L.require_cuda(w)
    co, ci = w.shape
    assert ci % group_size == 0
    wc = w.detach().contiguous()
    x = input_feat.view(-1, input_feat.shape[-1])
    assert x.shape[-1] == ci
    # token sub-sampling x[:, 0::n_tok // n_sample_token] (autoclip.py:27) becomes a row stride, not a copy
    step = x.shape[0] // n_sample_token          # ZeroDivisionError semantics of the reference: step == 0 -> slice error
    if step == 0:
        raise ValueError("slice step cannot be zero")
    x = x.to(device=w.device, dtype=w.dtype)     # the reference moves the (CPU-staged) activations to the weight's device
    if x.stride(-1) != 1:
        x = x.contiguous()
    n_tok = (x.shape[0] + step - 1) // step
    row_stride = x.stride(0) * step
    ng = ci // group_size
    best_max = torch.empty((co, ng, 1), dtype=w.dtype, device=w.device)
    best_min = torch.empty((co, ng, 1), dtype=w.dtype, device=w.device)
    idx = torch.empty((co, ng), dtype=torch.int32, device=w.device) if return_info else None
    err = torch.empty((co, ng), dtype=torch.float32, device=w.device) if return_info else None
    with torch.cuda.device(w.device):
        rc = L.lib().bdq_autoclip(L.ptr(wc), L.ptr(x), L.ptr(best_max), L.ptr(best_min), L.ptr(idx), L.ptr(err),
                                  L.i64(co), L.i64(ci), L.i64(n_tok), L.i64(row_stride), int(group_size), int(n_bit),
                                  L.QUANT_INT if quant_type == "int" else L.QUANT_NF3, int(n_grid),
                                  L.ctypes.c_double(max_shrink), L.dtype_code(wc), L.semantics(), L.stream_ptr(w.device))
    L.check(rc, "bdq_autoclip")
    if return_info:
        return best_max, best_min, idx, err
    return best_max, best_min


auto_clip_layer = auto_2clip_layer  # the name BASELINE.json's north_star uses


@torch.no_grad()
def auto_clip_block(module, w_bit, q_config, input_feat):
    """autoclip.py:86-107: every nn.Linear of the block except q/k projections."""
    named_linears = {name: m for name, m in module.named_modules() if isinstance(m, nn.Linear)}
    clip_list = []
    for name in named_linears:
        # due to qk bmm, it is hard to clip precisely
        if any([_ in name for _ in ["q_", "k_", "query", "key", "Wqkv"]]):
            continue
        named_linears[name].cuda()
        max_val, min_val = auto_2clip_layer(named_linears[name].weight, input_feat[name], n_bit=w_bit, q_config=q_config)
        clip_list.append((name, max_val, min_val))
        named_linears[name].cpu()
    gc.collect()
    return clip_list
