"""Drop-in for the hot-path part of the reference's quantization/autoclip.py: auto_2clip_layer (autoclip.py:13-83) and
auto_clip_block (:86-107).  One kernel launch (bdq_autoclip) evaluates every candidate of every (output channel, group)
pair; the reference loops 100 times over a [oc_batch, 512, ci] broadcast product.  run_clip (autoclip.py:109-207) is the
layer loop with the activations captured and kept ON the GPU (the reference stages every linear's input
on the host, autoclip.py:170-173, and moves each linear to the GPU and back, :98-105); the CLI and the hub dataset loaders
are out of scope."""
import functools
import gc
from collections import defaultdict

import torch
import torch.nn as nn

from . import _lib as L
from .clip_utils import (append_str_prefix, apply_clip, get_blocks, get_calib_dataset, get_named_linears,  # noqa: F401
                         get_op_name)  # (autoclip.py does `from clip_utils import *`)
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
    with L.on(w.device):
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


_SKIP = ["q_", "k_", "query", "key", "Wqkv"]


@torch.no_grad()
def auto_clip_block_resident(module, w_bit, q_config, input_feat):
    """auto_clip_block for a block that already lives on the GPU with GPU-resident activations: no .cuda()/.cpu() per
    linear, no host copies.  Same result list as auto_clip_block."""
    clip_list = []
    for name, lin in get_named_linears(module).items():
        if any([_ in name for _ in _SKIP]):
            continue
        max_val, min_val = auto_2clip_layer(lin.weight, input_feat[name], n_bit=w_bit, q_config=q_config)
        clip_list.append((name, max_val, min_val))
    return clip_list


@torch.no_grad()
def run_clip(model, enc, w_bit, q_config, n_samples=128, seqlen=1024, datasets="pile", samples=None, device=None,
             sharded=False):
    """autoclip.py:109-207 with the whole loop resident on one GPU: the model is moved once (a 7B bf16 model plus the
    [n_blocks, seqlen, hidden] activations of one layer is ~20 GB of the 180 GB), the forward hooks keep REFERENCES to the
    linears' inputs instead of host copies, and bdq_autoclip reads the 512 sampled tokens through a row stride.  Returns the
    reference's {"clip": [(global name, max_val, min_val), ...]} and leaves the clipped weights in the model
    (autoclip.py:197).  `samples`: token blocks [n_blocks, seqlen] (or the list of [1, seqlen] blocks get_calib_dataset
    returns); if None they come from get_calib_dataset(datasets, enc, n_samples, seqlen).
    sharded=True under torch.distributed: layer i's input does not depend on the clipping of layers < i (the reference
    forwards the block BEFORE it clips it, autoclip.py:181-197), so every rank runs the cheap forward of all blocks, searches
    only the layers it owns and one all-gather completes the list (sharded.gather_clip_cache); the weights of all layers
    are then clipped from the gathered list, so every rank ends with the same model."""
    if samples is None:
        print(f"Using {datasets} dataset to do calibation")
        samples = get_calib_dataset(datasets=datasets, tokenizer=enc, n_samples=n_samples, block_size=seqlen)
    if isinstance(samples, (list, tuple)):
        samples = torch.cat(list(samples), dim=0)
    dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
    layers = get_blocks(model)
    model.to(dev)

    inps, layer_kwargs = [], {}

    class Catcher(nn.Module):
        def __init__(self, module):
            super().__init__()
            self.module = module

        def __getattr__(self, name):            # newer transformers read attributes (attention_type) off the layer
            try:
                return super().__getattr__(name)
            except AttributeError:
                return getattr(self.module, name)

        def forward(self, inp, **kwargs):
            inps.append(inp)
            layer_kwargs.update(kwargs)
            raise ValueError  # early exit to break later inference

    layers[0] = Catcher(layers[0])
    try:
        try:
            # no KV cache: the captured layer kwargs would otherwise carry a cache that every block forward appends to
            # (2 GB per layer at 128 x 1024 tokens of a 7B model)
            model(samples.to(dev), use_cache=False)
        except TypeError:
            model(samples.to(dev))
    except ValueError:
        pass
    layers[0] = layers[0].module
    inps = inps[0]

    if sharded:
        import torch.distributed as dist

        from .sharded import gather_clip_cache, owned_layers
        mine = set(owned_layers(len(layers), dist.get_rank(), dist.get_world_size()))
    else:
        mine = set(range(len(layers)))

    clip_results = {"clip": []}
    for i in range(len(layers)):
        layer = layers[i]
        named_linears = get_named_linears(layer)
        input_feat = defaultdict(list)
        handles = []
        if i in mine:
            def cache_input_hook(m, x, y, name, feat_dict):
                feat_dict[name].append(x[0].detach())     # stays on the GPU; q/k/v and gate/up share one tensor
            for name in named_linears:
                if any([_ in name for _ in _SKIP]):
                    continue
                handles.append(named_linears[name].register_forward_hook(
                    functools.partial(cache_input_hook, name=name, feat_dict=input_feat)))
        out = layer(inps, **layer_kwargs)
        inps = out[0] if isinstance(out, (tuple, list)) else out
        for h in handles:
            h.remove()
        if i not in mine:
            continue
        input_feat = {k: (v[0] if len(v) == 1 else torch.cat(v, dim=0)) for k, v in input_feat.items()}
        clip_list = auto_clip_block_resident(layer, w_bit=w_bit, q_config=q_config, input_feat=input_feat)
        if not sharded:
            apply_clip(layer, clip_list)
        clip_results["clip"] += append_str_prefix(clip_list, get_op_name(model, layer) + ".")
        del input_feat
    if sharded:
        wdtype = next(layers[0].parameters()).dtype
        clip_results["clip"] = gather_clip_cache(clip_results["clip"], dtype=wdtype, device=dev)
        apply_clip(model, clip_results)
    return clip_results
