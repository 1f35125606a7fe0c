"""real_quantize_model_weight (quantization/quantizer.py:73-108) on a whole (tiny, random-init)
LlamaForCausalLM, the packed state dict as the on-disk format (inference/dump_quant.py:70-99 -> torch.save;
inference/utils/load_quant.py: init_only=True skeleton + load_state_dict), and make_fused_mlp on the loaded model."""
import io

import pytest
import torch

pytestmark = pytest.mark.gpu


def _tiny_llama(seed):
    from transformers import LlamaConfig, LlamaForCausalLM
    cfg = LlamaConfig(vocab_size=160, hidden_size=256, intermediate_size=768, num_hidden_layers=2, num_attention_heads=4,
                      num_key_value_heads=4, max_position_embeddings=64, tie_word_embeddings=False)
    torch.manual_seed(seed)
    return LlamaForCausalLM(cfg).half().eval()


@torch.no_grad()
def test_real_quantize_export_load_roundtrip():
    from bitdistiller_b200.fused_mlp import QuantLlamaMLP, make_fused_mlp
    from bitdistiller_b200.qmodule import WQLinear
    from bitdistiller_b200.quantizer import pseudo_quantize_tensor, real_quantize_model_weight
    dev = torch.device("cuda:0")
    q_config = {"zero_point": True, "q_group_size": 128}
    ids1 = torch.tensor([[7]], device=dev)
    ids12 = torch.arange(3, 15, device=dev).view(1, 12)

    # fake-quantised fp16 model: the numerical reference (same dequantised weights, dense matmuls)
    fake = _tiny_llama(0).to(dev)
    for layer in fake.model.layers:
        for m in layer.modules():
            if isinstance(m, torch.nn.Linear):
                m.weight.data = pseudo_quantize_tensor(m.weight.data, n_bit=2, **q_config)
    ref1, ref12 = fake(ids1).logits.float(), fake(ids12).logits.float()

    # packed model: every decoder-block linear becomes a WQLinear, nothing else changes
    # (the reference's flow, dump_quant.py:51-58: model on the host, every linear visits the GPU, whole model moves after)
    real = _tiny_llama(0)
    real_quantize_model_weight(real, 2, q_config)
    real = real.to(dev)
    n_lin = sum(isinstance(m, torch.nn.Linear) for l in real.model.layers for m in l.modules())
    n_wq = sum(isinstance(m, WQLinear) for l in real.model.layers for m in l.modules())
    assert n_lin == 0 and n_wq == 2 * 7 and isinstance(real.lm_head, torch.nn.Linear)
    out1, out12 = real(ids1).logits.float(), real(ids12).logits.float()    # 1 row: GEMV, 12 rows: tcgen05 GEMM
    assert float((out1 - ref1).norm() / ref1.norm()) < 2e-2
    assert float((out12 - ref12).norm() / ref12.norm()) < 2e-2

    # on-disk format = the state dict: qweight int32 [out, in/16], qzeros int32 [out, zw], scales fp16 [out, 16*zw]
    sd = real.state_dict()
    k = "model.layers.1.mlp.down_proj."
    assert sd[k + "qweight"].dtype == torch.int32 and tuple(sd[k + "qweight"].shape) == (256, 768 // 16)
    assert sd[k + "qzeros"].dtype == torch.int32 and tuple(sd[k + "qzeros"].shape) == (256, 1)
    assert sd[k + "scales"].dtype == torch.float16 and tuple(sd[k + "scales"].shape) == (256, 16)
    buf = io.BytesIO()
    torch.save(sd, buf)
    buf.seek(0)
    loaded_sd = torch.load(buf, map_location="cpu")

    # loader: skeleton with init_only=True (no quantisation work), then load_state_dict (strict)
    skel = _tiny_llama(1)                       # different random weights: everything must come from the file
    real_quantize_model_weight(skel, 2, q_config, init_only=True)
    skel.load_state_dict(loaded_sd, strict=True)
    skel = skel.to(dev)
    assert torch.equal(skel(ids1).logits, real(ids1).logits)
    assert torch.equal(skel(ids12).logits, real(ids12).logits)

    # fused MLP on the loaded model (fused_mlp.py:117-133): same packed buffers, decode through the fused launch
    make_fused_mlp(skel)
    assert all(isinstance(l.mlp, QuantLlamaMLP) for l in skel.model.layers)
    f1, f12 = skel(ids1).logits.float(), skel(ids12).logits.float()
    assert float((f1 - out1).norm() / out1.norm()) < 5e-3
    assert float((f12 - out12).norm() / out12.norm()) < 5e-3


@torch.no_grad()
def test_fused_qkv_linear_equals_separate():
    """fused_attn.py:312-337: one WQLinear over the concatenated packed buffers == the three projections, decode and prefill."""
    from bitdistiller_b200.fused_attn import fuse_qkv_linear, split_qkv
    from bitdistiller_b200.quantizer import real_quantize_model_weight
    dev = torch.device("cuda:0")
    m = _tiny_llama(2)
    real_quantize_model_weight(m, 2, {"zero_point": True, "q_group_size": 128})
    att = m.model.layers[0].self_attn.to(dev)
    qkv = fuse_qkv_linear(att.q_proj, att.k_proj, att.v_proj)
    assert qkv.out_features == 3 * 256 and tuple(qkv.qweight.shape) == (768, 16)
    for rows in (1, 5, 40):
        x = torch.randn(1, rows, 256, device=dev, dtype=torch.float16)
        q, k, v = split_qkv(qkv(x), 256, 256, 256)
        for got, proj in ((q, att.q_proj), (k, att.k_proj), (v, att.v_proj)):
            ref = proj(x)
            assert float((got.float() - ref.float()).norm() / ref.float().norm()) < 2e-3, rows


@torch.no_grad()
def test_make_quant_attn_fused_qkv_on_packed_model():
    """SURVEY 8f N2, second half: make_quant_attn (inference/modules/fused_attn.py:304-340) on a packed model -- one fused
    q/k/v projection per attention block, same logits as the three separate packed projections (decode and prefill)."""
    from bitdistiller_b200 import _lib as L
    from bitdistiller_b200.fused_attn import QuantLlamaAttention, make_quant_attn
    from bitdistiller_b200.quantizer import real_quantize_model_weight
    dev = torch.device("cuda:0")
    q_config = {"zero_point": True, "q_group_size": 128}
    ids1 = torch.tensor([[7]], device=dev)
    ids12 = torch.arange(3, 15, device=dev).view(1, 12)
    real = _tiny_llama(0)
    real_quantize_model_weight(real, 2, q_config)
    real = real.to(dev)
    ref1, ref12 = real(ids1).logits.float(), real(ids12).logits.float()
    n0 = L.launch_count()
    real(ids1)
    per_token_before = L.launch_count() - n0
    make_quant_attn(real, dev)
    assert sum(isinstance(m, QuantLlamaAttention) for m in real.modules()) == 2
    out1, out12 = real(ids1).logits.float(), real(ids12).logits.float()
    assert float((out1 - ref1).norm() / ref1.norm()) < 5e-3
    assert float((out12 - ref12).norm() / ref12.norm()) < 5e-3
    n0 = L.launch_count()
    real(ids1)
    assert L.launch_count() - n0 < per_token_before               # q, k, v: three launches -> one, per layer
    # decode with a KV cache through generate(): same greedy tokens as the unfused packed model
    unf = _tiny_llama(0)
    real_quantize_model_weight(unf, 2, q_config)
    unf = unf.to(dev)
    a = real.generate(ids12, max_new_tokens=6, do_sample=False)
    b = unf.generate(ids12, max_new_tokens=6, do_sample=False)
    assert a.shape == b.shape == (1, 18)
