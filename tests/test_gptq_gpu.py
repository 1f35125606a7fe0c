"""GPTQ-v2 2-bit export (train/convert_hf_to_gptq_manual.py:11-157) through libbdq.so.
Bit-exact against the reference's own outputs (tests/golden/gptq.npz), against the oracle on fresh inputs, and through
unpack -> dequant round trips at Llama-2-7B sizes."""
import numpy as np
import pytest
import torch

import oracle
from _util import DT, bits_equal, load_cases, t

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _linear(w, bias=False):
    lin = torch.nn.Linear(w.shape[1], w.shape[0], bias=bias, dtype=w.dtype)
    lin.weight.data = w
    return lin


def _unpack(qweight, qzeros, scales, G):
    """GPTQ-v2 tensors -> (codes [OC, IC], zeros [OC, IC/G], dequantised fp32 [OC, IC])."""
    sh = torch.arange(16, device=qweight.device, dtype=torch.int32) * 2
    codes = ((qweight.unsqueeze(1) >> sh.view(1, 16, 1)) & 3).reshape(-1, qweight.shape[1]).t()          # [OC, IC]
    zeros = ((qzeros.unsqueeze(2) >> sh.view(1, 1, 16)) & 3).reshape(qzeros.shape[0], -1).t()            # [OC, IC/G]
    s = scales.t().float()
    dq = (codes.float().view(codes.shape[0], -1, G) - zeros.float().unsqueeze(-1)) * s.unsqueeze(-1)
    return codes.contiguous(), zeros.contiguous(), dq.reshape(codes.shape)


def test_fused_export_matches_reference_golden():
    from bitdistiller_b200.convert_hf_to_gptq_manual import convert_linear_to_gptq
    n = 0
    for meta, c in load_cases("gptq"):
        dname, G = meta[0], int(meta[1])
        for dev in (DEV, "cpu"):   # a host-resident module (the reference's flow) converts through the GPU as well
            lin = _linear(t(c["w"], DT[dname])).to(dev)
            qw, qz, sc, g_idx, bias = convert_linear_to_gptq(lin, 2, G)
            assert qw.device.type == torch.device(dev).type and qw.dtype == torch.int32 and sc.dtype == torch.bfloat16
            assert np.array_equal(qw.cpu().numpy(), c["qweight"]), meta
            assert np.array_equal(qz.cpu().numpy(), c["qzeros"]), meta
            assert bits_equal(sc.cpu(), t(c["scales"], torch.bfloat16)), meta
            assert np.array_equal(g_idx.cpu().numpy(), c["g_idx"]), meta
            assert bias is None
        n += 1
    assert n == 24


def test_two_step_api_matches_reference_golden():
    from bitdistiller_b200.convert_hf_to_gptq_manual import convert_layer_to_gptq_format, get_hf_quant_params
    for meta, c in load_cases("gptq"):
        dname, G = meta[0], int(meta[1])
        dt = DT[dname]
        lin = _linear(t(c["w"], dt), bias=True).to(DEV)
        codes, s_t, z_t, bias = get_hf_quant_params(lin, 2, True, G)
        assert codes.dtype == dt and s_t.dtype == dt and z_t.dtype == dt and bias is lin.bias
        assert s_t.shape == (lin.in_features // G, lin.out_features)
        assert bits_equal(codes.cpu(), t(c["codes"], dt)), meta
        assert bits_equal(s_t.cpu(), t(c["s_t"], dt)), meta
        assert bits_equal(z_t.cpu(), t(c["z_t"], dt)), meta
        qw, qz, sc, g_idx, b = convert_layer_to_gptq_format(codes, s_t, z_t, bias, 2, G)
        assert np.array_equal(qw.cpu().numpy(), c["qweight"]), meta
        assert np.array_equal(qz.cpu().numpy(), c["qzeros"]), meta
        assert bits_equal(sc.cpu(), t(c["scales"], torch.bfloat16)), meta
        assert np.array_equal(g_idx.cpu().numpy(), c["g_idx"]), meta
        assert torch.equal(b, lin.bias.data) and b.data_ptr() != lin.bias.data_ptr()


@pytest.mark.parametrize("dname", ["bf16", "f16", "f32"])
@pytest.mark.parametrize("oc,ic,G", [(48, 1536, 128), (16, 11008, 64), (80, 4096, 128), (32, 544, 32), (16, 2048, 512)])
def test_fused_export_matches_oracle(dname, oc, ic, G):
    """Ragged tiles: OC not a multiple of 32, IC not a multiple of 512."""
    from bitdistiller_b200.convert_hf_to_gptq_manual import convert_linear_to_gptq
    g = torch.Generator().manual_seed(oc * 7 + ic + G)
    w = (torch.randn(oc, ic, generator=g) ** 3 * 0.02).to(DT[dname])
    w[0, :G] = 0.0                         # a constant group: s = 1e-5 / 3
    w[1, :G] = w[1, :G].abs()              # min >= 0: z clamps to 0
    w[2, :G] = -w[2, :G].abs() - 0.01      # max < 0: z clamps to 3
    qw, qz, sc, _, _ = convert_linear_to_gptq(_linear(w.clone()).to(DEV), 2, G)
    oqw, oqz, osc = oracle.gptq_export_w2(w, G)
    assert torch.equal(qw.cpu(), oqw)
    assert torch.equal(qz.cpu(), oqz)
    assert bits_equal(sc.cpu(), osc)


@pytest.mark.parametrize("oc,ic", [(4096, 4096), (11008, 4096), (4096, 11008)])
def test_full_size_round_trip(oc, ic):
    """Llama-2-7B shapes: unpacked codes/zeros equal the fp32 fake-quantiser's, a row slice equals the oracle, and the
    two-step API packs to the same bytes as the fused kernel."""
    from bitdistiller_b200.convert_hf_to_gptq_manual import (convert_layer_to_gptq_format, convert_linear_to_gptq,
                                                             get_hf_quant_params)
    G = 128
    torch.manual_seed(oc + ic)
    w = (torch.randn(oc, ic, device=DEV) * 0.02).to(torch.bfloat16)
    lin = _linear(w)
    qw, qz, sc, g_idx, _ = convert_linear_to_gptq(lin, 2, G)
    assert qw.shape == (ic // 16, oc) and qz.shape == (ic // G, oc // 16) and sc.shape == (ic // G, oc)
    assert torch.equal(g_idx, torch.arange(ic, device=DEV, dtype=torch.int32) // G)
    codes, zeros, _ = _unpack(qw, qz, sc, G)
    c2, s_t, z_t, _ = get_hf_quant_params(lin, 2, True, G)
    assert torch.equal(codes, c2.to(torch.int32))
    assert torch.equal(zeros, z_t.t().to(torch.int32))
    assert torch.equal(sc, s_t)
    qw2, qz2, sc2, _, _ = convert_layer_to_gptq_format(c2, s_t, z_t, None, 2, G)
    assert torch.equal(qw, qw2) and torch.equal(qz, qz2) and torch.equal(sc, sc2)
    rows = torch.arange(0, oc, oc // 16)[:16]
    oqw, oqz, osc = oracle.gptq_export_w2(w[rows.to(DEV)].cpu(), G)
    assert torch.equal(qw[:, rows.to(DEV)].cpu(), oqw)
    assert bits_equal(sc[:, rows.to(DEV)].cpu(), osc)


def test_dequantised_export_equals_fake_quantiser():
    """(code - zero) * scale of the export, with the fp32 scale, is pseudo_quantize_tensor of the fp32 weight."""
    from bitdistiller_b200.convert_hf_to_gptq_manual import convert_linear_to_gptq
    from bitdistiller_b200.quantizer import pseudo_quantize_tensor
    G = 64
    torch.manual_seed(3)
    w = torch.randn(256, 1024, device=DEV) * 0.02
    qw, qz, sc, _, _ = convert_linear_to_gptq(_linear(w), 2, G)
    codes, zeros, _ = _unpack(qw, qz, sc, G)
    dq, s, z = pseudo_quantize_tensor(w, n_bit=2, zero_point=True, q_group_size=G, get_scale_zp=True)
    assert torch.equal(zeros.float(), z.view(256, -1))
    assert torch.equal(sc.t(), s.view(256, -1).to(torch.bfloat16))
    mine = (codes.float().view(256, -1, G) - z.view(256, -1, 1)) * s.view(256, -1, 1)
    assert torch.equal(mine.view(256, 1024), dq)


def test_error_behaviour_matches_reference():
    from bitdistiller_b200 import _lib
    from bitdistiller_b200.convert_hf_to_gptq_manual import (convert_hf_to_gptq, convert_layer_to_gptq_format,
                                                             convert_linear_to_gptq)
    lin = _linear(torch.randn(32, 256, device=DEV).to(torch.bfloat16))
    with pytest.raises(ValueError, match="bits=2"):                       # ref 89-90
        convert_linear_to_gptq(lin, 4, 64)
    with pytest.raises(ValueError, match="2-bit"):                        # ref 172-173
        convert_hf_to_gptq("/nonexistent", "/nonexistent", bits=3)
    with pytest.raises(ValueError, match="out_features"):                 # ref 123-124
        convert_linear_to_gptq(_linear(torch.randn(24, 256, device=DEV)), 2, 64)
    with pytest.raises(ValueError, match="in_features"):                  # ref 107-108
        convert_linear_to_gptq(_linear(torch.randn(32, 40, device=DEV)), 2, 8)
    codes = torch.randint(0, 4, (32, 256), device=DEV).to(torch.bfloat16)
    bad_z = torch.full((4, 32), 4.0, device=DEV, dtype=torch.bfloat16)
    with pytest.raises(ValueError, match="Zero points"):                  # ref 103-104
        convert_layer_to_gptq_format(codes, torch.ones(4, 32, device=DEV), bad_z, None, 2, 64)
    with pytest.raises(_lib.BdqError, match="group=48"):                  # not a power of two: refused, never wrong
        convert_linear_to_gptq(_linear(torch.randn(32, 96, device=DEV)), 2, 48)


def test_state_dict_conversion():
    """convert_hf_to_gptq's layer loop (ref 196-237) on a small host-resident model."""
    from bitdistiller_b200.convert_hf_to_gptq_manual import (convert_linear_to_gptq, convert_state_dict_to_gptq,
                                                             gptq_quantize_config)

    class Block(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.q_proj = torch.nn.Linear(128, 128, bias=True)
            self.down_proj = torch.nn.Linear(256, 128, bias=False)
            self.norm = torch.nn.LayerNorm(128)
            self.lm_head = torch.nn.Linear(128, 64, bias=False)

    torch.manual_seed(5)
    model = torch.nn.ModuleDict({"layer0": Block(), "layer1": Block()}).to(torch.bfloat16)
    sd, names = convert_state_dict_to_gptq(model, 2, 64)
    assert names == ["layer0.q_proj", "layer0.down_proj", "layer1.q_proj", "layer1.down_proj"]
    for n in names:
        assert f"{n}.weight" not in sd
        qw, qz, sc, g_idx, bias = convert_linear_to_gptq(model.get_submodule(n), 2, 64)
        assert torch.equal(sd[f"{n}.qweight"], qw) and torch.equal(sd[f"{n}.qzeros"], qz)
        assert torch.equal(sd[f"{n}.scales"], sc) and torch.equal(sd[f"{n}.g_idx"], g_idx)
        assert (f"{n}.bias" in sd) == (bias is not None)
        assert all(not sd[f"{n}.{k}"].is_cuda for k in ("qweight", "qzeros", "scales", "g_idx"))
    assert "layer0.lm_head.weight" in sd and "layer0.norm.weight" in sd
    cfg = gptq_quantize_config(2, 64)
    assert cfg["checkpoint_format"] == "gptq_v2" and cfg["bits"] == 2 and cfg["sym"] is False and cfg["desc_act"] is False
