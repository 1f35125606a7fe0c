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


class QuantLlamaAttention(torch.nn.Module):
    """The module `make_quant_attn` installs (fused_attn.py:213-302 in the reference): ONE packed projection for q, k and v,
    the packed o_proj, and the attention math left to the installed transformers version (its rotary helper, KV cache and
    attention interface are called exactly as `LlamaAttention.forward` calls them) -- only the projections are this
    package's kernels: one GEMV launch (decode) or one tcgen05 GEMM (prefill) instead of three."""

    def __init__(self, attn, qkv_proj):
        super().__init__()
        self.config, self.layer_idx = attn.config, attn.layer_idx
        self.head_dim, self.scaling = attn.head_dim, attn.scaling
        self.num_key_value_groups = attn.num_key_value_groups
        self.attention_dropout, self.is_causal = attn.attention_dropout, getattr(attn, "is_causal", True)
        self.q_out, self.k_out, self.v_out = attn.q_proj.out_features, attn.k_proj.out_features, attn.v_proj.out_features
        self.qkv_proj, self.o_proj = qkv_proj, attn.o_proj

    def forward(self, hidden_states, position_embeddings=None, attention_mask=None, past_key_values=None, **kwargs):
        from transformers.models.llama.modeling_llama import (ALL_ATTENTION_FUNCTIONS, apply_rotary_pos_emb,
                                                              eager_attention_forward)
        input_shape = hidden_states.shape[:-1]
        hidden_shape = (*input_shape, -1, self.head_dim)
        q, k, v = split_qkv(self.qkv_proj(hidden_states), self.q_out, self.k_out, self.v_out)
        q = q.reshape(hidden_shape).transpose(1, 2)
        k = k.reshape(hidden_shape).transpose(1, 2)
        v = v.reshape(hidden_shape).transpose(1, 2)
        cos, sin = position_embeddings
        q, k = apply_rotary_pos_emb(q, k, cos, sin)
        if past_key_values is not None:
            k, v = past_key_values.update(k, v, self.layer_idx)
        attention_interface = ALL_ATTENTION_FUNCTIONS.get_interface(self.config._attn_implementation, eager_attention_forward)
        out, weights = attention_interface(self, q, k, v, attention_mask, dropout=0.0 if not self.training else self.attention_dropout,
                                           scaling=self.scaling, **kwargs)
        out = out.reshape(*input_shape, -1).contiguous()
        return self.o_proj(out), weights


def make_quant_attn(model, dev=None):
    """fused_attn.py:304-340: replace every LlamaAttention whose projections are packed WQLinears by QuantLlamaAttention
    with the fused q/k/v projection.  Unlike the reference the model is not moved to the host and back."""
    from .pre_quant import set_op_by_name
    for name, m in list(model.named_modules()):
        if m.__class__.__name__ not in ("LlamaAttention", "LlamaAttentionFused"):
            continue
        if not isinstance(m.q_proj, WQLinear):
            continue
        attn = QuantLlamaAttention(m, fuse_qkv_linear(m.q_proj, m.k_proj, m.v_proj))
        if dev is not None:
            attn.to(dev)
        set_op_by_name(model, name, attn)
    return model
