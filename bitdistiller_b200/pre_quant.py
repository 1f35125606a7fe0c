"""The three helpers of the reference's quantization/pre_quant.py that the hot path uses (pre_quant.py:16-49, called
from quantizer.py:78-79).  The rest of that file is dead AWQ code and is out of scope."""
import torch.nn as nn


def get_named_linears(module):
    return {name: m for name, m in module.named_modules() if isinstance(m, nn.Linear)}


def set_op_by_name(layer, name, new_module):
    levels = name.split('.')
    if len(levels) > 1:
        mod_ = layer
        for l_idx in range(len(levels) - 1):
            if levels[l_idx].isdigit():
                mod_ = mod_[int(levels[l_idx])]
            else:
                mod_ = getattr(mod_, levels[l_idx])
        setattr(mod_, levels[-1], new_module)
    else:
        setattr(layer, name, new_module)


def get_blocks(model):
    """Decoder-block list per model family (pre_quant.py:31-49); resolved by class name so that no model code is imported."""
    cname = model.__class__.__name__
    low = str(model.__class__).lower()
    if cname in ("LlamaForCausalLM", "Qwen3ForCausalLM", "Qwen2ForCausalLM", "MistralForCausalLM"):
        return model.model.layers
    if cname == "OPTForCausalLM":
        return model.model.decoder.layers
    if cname == "BloomForCausalLM":
        return model.transformer.h
    if "mpt" in low:
        return model.transformer.blocks
    if "falcon" in low or "bigcode" in low:
        return model.transformer.h
    if "neox" in low:
        return model.gpt_neox.layers
    raise NotImplementedError(type(model))
