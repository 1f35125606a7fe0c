"""The weight-side half of the reference's fused attention (inference/modules/fused_attn.py:304-340, `make_quant_attn`):
q, k and v become ONE packed WQLinear by concatenating their qweight / qzeros / scales (and bias) along out_features,
so a decode step issues one GEMV and a prefill one tcgen05 GEMM for the three projections.  The attention math itself
(rotary embedding, KV cache, softmax) is outside the hot path this package covers (out of scope)."""
import torch

from .qmodule import WQLinear


def fuse_qkv_linear(q_proj, k_proj, v_proj):
    """fused_attn.py:312-337: returns a WQLinear with out_features = q + k + v."""
    for p in (k_proj, v_proj):
        if (p.w_bit, p.group_size, p.in_features) != (q_proj.w_bit, q_proj.group_size, q_proj.in_features):
            raise ValueError("q/k/v projections must share w_bit, group_size and in_features")
    has_bias = q_proj.bias is not None
    qkv = WQLinear(q_proj.w_bit, q_proj.group_size, q_proj.in_features,
                   q_proj.out_features + k_proj.out_features + v_proj.out_features, has_bias, q_proj.qweight.device)
    qkv.qweight = torch.cat([q_proj.qweight, k_proj.qweight, v_proj.qweight], dim=0).contiguous()
    qkv.qzeros = torch.cat([q_proj.qzeros, k_proj.qzeros, v_proj.qzeros], dim=0).contiguous()
    qkv.scales = torch.cat([q_proj.scales, k_proj.scales, v_proj.scales], dim=0).contiguous()
    qkv.bias = torch.cat([q_proj.bias, k_proj.bias, v_proj.bias], dim=0) if has_bias else None
    qkv.split_k_iters = q_proj.split_k_iters
    return qkv


def split_qkv(y, q_out, k_out, v_out):
    """[..., q + k + v] -> (q, k, v) views, the `torch.split` of fused_attn.py's forward."""
    return torch.split(y, [q_out, k_out, v_out], dim=-1)
