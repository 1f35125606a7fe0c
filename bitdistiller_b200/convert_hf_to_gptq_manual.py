"""GPTQ-v2 2-bit export of a QAT checkpoint: drop-in for the reference's `train/convert_hf_to_gptq_manual.py`.

Same function names, arguments and returned layouts; the per-layer arithmetic runs in libbdq.so:

  get_hf_quant_params            (ref 11-75)  -> bdq_int_fakequant_fwd on the fp32 weight (codes, scales, zeros)
  convert_layer_to_gptq_format   (ref 85-157) -> bdq_gptq_pack_w2 (K-major qweight, zeros packed along OC)
  convert_linear_to_gptq         (new)        -> bdq_gptq_export_w2: both steps in ONE pass over the weight
  convert_state_dict_to_gptq     (new)        -> the layer loop of convert_hf_to_gptq (ref 196-237) on a model object
  convert_hf_to_gptq             (ref 160-299)-> load / convert / save shell around it

The reference runs on a host-resident model (`model.cpu()`, ref 193); here each weight is copied to the GPU for the
kernel and the results are returned on the weight's own device, so a host model converts exactly as before.  The
arithmetic is always the reference's fp32 HOST arithmetic (true division), independent of BDQ_SEMANTICS.
"""
import json
import os

import torch

from . import _lib as L

_PROJ = ("q_proj", "k_proj", "v_proj", "o_proj", "gate_proj", "up_proj", "down_proj")   # ref 200-202


def _device():
    if not torch.cuda.is_available():
        raise L.BdqError("convert_hf_to_gptq_manual needs a CUDA device (no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device())


def _to_gpu(t):
    return t if t.is_cuda else t.to(_device())


@torch.no_grad()
def get_hf_quant_params(module, n_bit=2, zero_point=True, q_group_size=64):
    """(qweight_unpacked [OC, IC], scales.t() [IC/G, OC], zero_points.t() [IC/G, OC], bias), all in the weight's dtype
    except bias (returned as is) -- ref 11-75.  Codes and zeros hold small integers."""
    if not zero_point:
        raise NotImplementedError("zero_point=False is never used by the reference (ref 48)")
    w0 = module.weight
    org, dev = w0.dtype, w0.device
    oc, ic = w0.shape
    if q_group_size == -1:
        q_group_size = ic
    assert ic % q_group_size == 0
    w = _to_gpu(w0.detach()).to(torch.float32).contiguous()
    n, ng = w.numel(), w.numel() // q_group_size
    dq = torch.empty_like(w)
    scales = torch.empty(ng, dtype=torch.float32, device=w.device)
    zeros = torch.empty(ng, dtype=torch.float32, device=w.device)
    codes = torch.empty(w.shape, dtype=torch.uint8, device=w.device)
    with L.on(w.device):
        rc = L.lib().bdq_int_fakequant_fwd(L.ptr(w), L.ptr(dq), L.ptr(scales), L.ptr(zeros), L.ptr(codes), L.ptr(None),
                                           L.ptr(None), L.i64(n), int(q_group_size), int(n_bit), L.F32,
                                           L.ROUND_HALF_EVEN, L.SEM_CPU, L.stream_ptr(w.device))
    L.check(rc, "bdq_int_fakequant_fwd")
    if os.environ.get("BDQ_CHECK_NAN", "0") == "1":
        assert torch.isnan(scales).sum() == 0
        assert torch.isnan(w).sum() == 0
    out = lambda t: t.to(org).to(dev)
    return (out(codes), out(scales.view(oc, -1)).t(), out(zeros.view(oc, -1).to(torch.int32)).t(), module.bias)


@torch.no_grad()
def convert_layer_to_gptq_format(hf_qweight_unpacked, hf_scales, hf_zero_points_unpacked, bias_orig, bits, group_size):
    """Unpacked parameters (as returned by get_hf_quant_params) -> (qweight int32 [IC/16, OC], qzeros int32
    [IC/G, OC/16], scales bf16 [IC/G, OC], g_idx int32 [IC], bias) -- ref 85-157."""
    if bits != 2:
        raise ValueError("This conversion script is designed for bits=2.")
    oc, ic = hf_qweight_unpacked.shape
    dev = hf_qweight_unpacked.device
    if group_size == -1:
        raise ValueError("group_size=-1 is not supported (the reference computes a negative group count, ref 121)")
    zp = hf_zero_points_unpacked.t()                       # [OC, IC/G]
    if not ((zp >= 0) & (zp < (1 << bits))).all():
        raise ValueError(f"Zero points must be in the range [0, {(1 << bits) - 1}] for {bits}-bit quantization.")
    if ic % 16 != 0:
        raise ValueError(f"in_features ({ic}) must be divisible by pack_factor (16)")
    if oc % 16 != 0:
        raise ValueError(f"out_features ({oc}) must be divisible by pack_factor (16) for zero points")
    codes = _to_gpu(hf_qweight_unpacked).to(torch.uint8).contiguous()
    zeros = _to_gpu(zp).to(torch.uint8).contiguous()
    qweight = torch.empty((ic // 16, oc), dtype=torch.int32, device=codes.device)
    qzeros = torch.empty((ic // group_size, oc // 16), dtype=torch.int32, device=codes.device)
    with L.on(codes.device):
        rc = L.lib().bdq_gptq_pack_w2(L.ptr(codes), L.ptr(zeros), L.ptr(qweight), L.ptr(qzeros), L.i64(oc), L.i64(ic),
                                      int(group_size), int(bits), L.stream_ptr(codes.device))
    L.check(rc, "bdq_gptq_pack_w2")
    scales = hf_scales.contiguous().to(torch.bfloat16)
    g_idx = torch.arange(ic, dtype=torch.int32, device=dev) // group_size
    bias = bias_orig.data.clone() if bias_orig is not None else None
    return qweight.to(dev), qzeros.to(dev), scales, g_idx, bias


@torch.no_grad()
def convert_linear_to_gptq(module, bits=2, group_size=64):
    """get_hf_quant_params + convert_layer_to_gptq_format fused: one kernel reads the weight once and writes the
    packed GPTQ tensors.  Returns (qweight, qzeros, scales, g_idx, bias) exactly like convert_layer_to_gptq_format."""
    if bits != 2:
        raise ValueError("This conversion script is designed for bits=2.")
    w0 = module.weight
    dev = w0.device
    oc, ic = w0.shape
    if group_size == -1:
        raise ValueError("group_size=-1 is not supported (the reference computes a negative group count, ref 121)")
    if ic % 16 != 0:
        raise ValueError(f"in_features ({ic}) must be divisible by pack_factor (16)")
    if oc % 16 != 0:
        raise ValueError(f"out_features ({oc}) must be divisible by pack_factor (16) for zero points")
    assert ic % group_size == 0
    w = _to_gpu(w0.detach()).contiguous()
    qweight = torch.empty((ic // 16, oc), dtype=torch.int32, device=w.device)
    qzeros = torch.empty((ic // group_size, oc // 16), dtype=torch.int32, device=w.device)
    scales = torch.empty((ic // group_size, oc), dtype=torch.bfloat16, device=w.device)
    with L.on(w.device):
        rc = L.lib().bdq_gptq_export_w2(L.ptr(w), L.ptr(qweight), L.ptr(qzeros), L.ptr(scales), L.i64(oc), L.i64(ic),
                                        int(group_size), int(bits), L.dtype_code(w), L.stream_ptr(w.device))
    L.check(rc, "bdq_gptq_export_w2")
    g_idx = torch.arange(ic, dtype=torch.int32, device=dev) // group_size
    bias = module.bias.data.clone() if module.bias is not None else None
    return qweight.to(dev), qzeros.to(dev), scales.to(dev), g_idx, bias


@torch.no_grad()
def convert_state_dict_to_gptq(model, bits=2, group_size=64):
    """The layer loop of convert_hf_to_gptq (ref 196-237): every q/k/v/o/gate/up/down nn.Linear loses `.weight` and
    gains `.qweight/.qzeros/.scales/.g_idx` (host tensors).  Returns (new_state_dict, converted_layer_names)."""
    new_state_dict = dict(model.state_dict())
    converted = []
    for name, module in model.named_modules():
        if isinstance(module, torch.nn.Linear) and any(p in name for p in _PROJ):
            qweight, qzeros, scales, g_idx, bias = convert_linear_to_gptq(module, bits, group_size)
            new_state_dict.pop(f"{name}.weight", None)
            if bias is not None:
                new_state_dict.pop(f"{name}.bias", None)
            new_state_dict[f"{name}.qweight"] = qweight.cpu()
            new_state_dict[f"{name}.qzeros"] = qzeros.cpu()
            new_state_dict[f"{name}.scales"] = scales.cpu()
            new_state_dict[f"{name}.g_idx"] = g_idx.cpu()
            if bias is not None:
                new_state_dict[f"{name}.bias"] = bias.cpu()
            converted.append(name)
    return new_state_dict, converted


def gptq_quantize_config(bits=2, group_size=64):
    """quantize_config.json of the export (ref 244-256)."""
    return {"bits": bits, "group_size": group_size, "sym": False, "desc_act": False, "model_file_base_name": "model",
            "quant_method": "gptq", "checkpoint_format": "gptq_v2", "meta": {"quantizer": "gptqmodel:2.2.0"}}


def convert_hf_to_gptq(hf_model_path, save_path, bits=2, group_size=64):
    """Load a Hugging Face checkpoint, convert its projection layers and save model.safetensors +
    quantize_config.json + config (+ tokenizer) under save_path -- ref 160-299."""
    if bits != 2:
        raise ValueError("This script is specialized for 2-bit conversion.")
    from safetensors.torch import save_file
    from transformers import AutoConfig, AutoModelForCausalLM
    config = AutoConfig.from_pretrained(hf_model_path, trust_remote_code=True)
    model = AutoModelForCausalLM.from_pretrained(hf_model_path, config=config, torch_dtype=torch.bfloat16,
                                                 trust_remote_code=True, low_cpu_mem_usage=True)
    model.eval()
    model.cpu()
    new_state_dict, converted = convert_state_dict_to_gptq(model, bits, group_size)
    if not converted:
        print("No layers were converted. Please check your model structure.")
        return
    quantize_config = gptq_quantize_config(bits, group_size)
    os.makedirs(save_path, exist_ok=True)
    with open(os.path.join(save_path, "quantize_config.json"), "w") as f:
        json.dump(quantize_config, f, indent=2)
    new_state_dict = {k: (v.cpu().contiguous() if isinstance(v, torch.Tensor) else v) for k, v in new_state_dict.items()}
    save_file(new_state_dict, os.path.join(save_path, "model.safetensors"), metadata={"format": "pt"})
    model.config.quantization_config = quantize_config
    model.config.save_pretrained(save_path)
    try:
        from transformers import AutoTokenizer
        AutoTokenizer.from_pretrained(hf_model_path, trust_remote_code=True).save_pretrained(save_path)
    except Exception as e:  # same tolerance as the reference (ref 291-296)
        print(f"Could not save tokenizer: {e}. You may need to copy it manually.")
    print(f"Converted layers: {converted}")
