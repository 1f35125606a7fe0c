"""Drop-in for the reference's fused MLP (inference/modules/fused_mlp.py:18-114): `QuantLlamaMLP` keeps the packed
gate / up buffers under the reference's names and `make_fused_mlp` swaps it in for every MLP whose three projections
are `WQLinear`s.  Decode (one row) runs silu(gate(x)) * up(x) as ONE launch of the batch-1 streaming GEMV with the
SiLU-and-multiply epilogue; 2..8 rows take the ring GEMV per projection, more rows the tcgen05 GEMM,
with the reference's roundings (fp16 after every op)."""
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import awq_inference_engine
from .qmodule import WQLinear


class QuantLlamaMLP(nn.Module):
    def __init__(self, gate_proj, down_proj, up_proj):
        super().__init__()
        self.register_buffer("gate_proj_qweight", gate_proj.qweight)
        self.register_buffer("gate_proj_scales", gate_proj.scales)
        self.register_buffer("gate_proj_qzeros", gate_proj.qzeros)
        self.register_buffer("up_proj_qweight", up_proj.qweight)
        self.register_buffer("up_proj_scales", up_proj.scales)
        self.register_buffer("up_proj_qzeros", up_proj.qzeros)
        self.in_features = gate_proj.in_features
        self.intermediate_size = gate_proj.out_features
        self.out_features = down_proj.out_features
        self.w_bit = gate_proj.w_bit
        self.down_proj = down_proj
        self.split_k_iters = down_proj.split_k_iters
        self.offset = 0x0F if self.w_bit == 4 else 0x03

    def forward(self, x):
        return self.down_proj(self.our_llama_mlp(x))

    @torch.no_grad()
    def our_llama_mlp(self, x):
        out_shape = x.shape[:-1] + (self.intermediate_size,)
        x = x.reshape(-1, x.shape[-1])
        G = self.down_proj.group_size
        gate = (self.gate_proj_qweight, self.gate_proj_scales, self.gate_proj_qzeros)
        up = (self.up_proj_qweight, self.up_proj_scales, self.up_proj_qzeros)
        if x.shape[0] == 1 and self.w_bit == 2 and G == 128:
            c = awq_inference_engine.gemv_silu_mul_forward_cuda(x, gate, up, G)
        elif x.shape[0] <= 8:
            g = F.silu(awq_inference_engine.gemv_forward_cuda(x, *gate, self.w_bit, G))
            c = g * awq_inference_engine.gemv_forward_cuda(x, *up, self.w_bit, G)
        else:
            g = F.silu(awq_inference_engine.gemm_forward_cuda(x, *gate, G, self.split_k_iters, w_bit=self.w_bit))
            c = g * awq_inference_engine.gemm_forward_cuda(x, *up, G, self.split_k_iters, w_bit=self.w_bit)
        return c.reshape(out_shape)


def make_fused_mlp(m, parent_name=""):
    """Replace every module that has WQLinear `gate_proj`, `up_proj` and `down_proj` children (LlamaMLP / Qwen3MLP
    layout; the reference matches on the LlamaMLP class, fused_mlp.py:117-133) by QuantLlamaMLP, in place."""
    for name, child in list(m.named_children()):
        projs = [getattr(child, k, None) for k in ("gate_proj", "down_proj", "up_proj")]
        if all(isinstance(p, WQLinear) for p in projs):
            setattr(m, name, QuantLlamaMLP(*projs))
        else:
            make_fused_mlp(child, parent_name=f"{parent_name}.{name}" if parent_name else name)
    return m
