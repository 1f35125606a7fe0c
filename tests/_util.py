"""Shared helpers for the parity tests."""
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
DT = {"f32": torch.float32, "bf16": torch.bfloat16, "f16": torch.float16}


def load_cases(name):
    """Yield (meta list[str], dict suffix->np.ndarray) for every case of tests/golden/<name>.npz."""
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    n = len([k for k in z.files if k.endswith("_meta") and k.startswith("c")])
    for i in range(n):
        pre = f"c{i}_"
        yield [str(s) for s in z[pre + "meta"]], {k[len(pre):]: z[k] for k in z.files if k.startswith(pre)}


def t(a, dtype):
    return torch.from_numpy(np.asarray(a)).to(dtype)


def bits_equal(a, b):
    """Value equality that treats NaN==NaN and +0==-0 (eager ops do not pin the sign of zero)."""
    a, b = a.to(torch.float64), b.to(torch.float64)
    return bool(((a == b) | (a.isnan() & b.isnan())).all())


def ulp_of(x, dtype):
    """Size of one ulp of dtype at |x| (elementwise, float64)."""
    x = x.to(torch.float64).abs()
    mant = {torch.float32: 23, torch.bfloat16: 7, torch.float16: 10}[dtype]
    tiny = {torch.float32: -126, torch.bfloat16: -126, torch.float16: -14}[dtype]
    e = torch.floor(torch.log2(torch.clamp(x, min=2.0 ** tiny)))
    return torch.pow(torch.tensor(2.0, dtype=torch.float64), e - mant)


# ---- plain eager-PyTorch restatements (run on whatever device the input lives on; used to confirm the
# ---- CPU-eager vs CUDA-eager rounding hypothesis of SURVEY.md App. D on the GPU box)
def torch_int_fakequant(w, n_bit, G, half_away):
    shp = w.shape
    x = w.reshape(-1, G)
    mx, mn = x.amax(1, keepdim=True), x.amin(1, keepdim=True)
    qmax = 2 ** n_bit - 1
    s = (mx - mn).clamp(min=1e-5) / qmax
    z = (-torch.round(mn / s)).clamp_(0, qmax)
    u = x / s
    r = torch.sign(u) * torch.floor(torch.abs(u) + 0.5) if half_away else torch.round(u)
    dq = (torch.clamp(r + z, 0, qmax) - z) * s
    return dq.reshape(shp), s.reshape(shp[0], -1), z.reshape(shp[0], -1)


def torch_nf3_fakequant(w, G):
    shp = w.shape
    x = w.reshape(-1, G)
    sp, sn = x.amax(1, keepdim=True).abs(), x.amin(1, keepdim=True).abs()
    zero = torch.zeros_like(x)
    qp = torch.where(x >= 0, x, zero) / sp
    qn = torch.where(x < 0, x, zero) / sn
    dev = x.device
    T = lambda v: torch.tensor(v).to(dev)
    yp = torch.where(qp >= 0.8114928305149078, T(1.0), qp)
    yp = torch.where((yp < 0.8114928305149078) & (yp >= 0.5024898052215576), T(0.6229856610298157), yp)
    yp = torch.where((yp < 0.5024898052215576) & (yp >= 0.2826657369732857), T(0.3819939494132996), yp)
    yp = torch.where((yp < 0.2826657369732857) & (yp >= 0.0916687622666359), T(0.1833375245332718), yp)
    yp = torch.where(yp < 0.0916687622666359, T(0), yp)
    yn = torch.where(qn >= -0.1234657019376755, T(0), qn)
    yn = torch.where((yn < -0.1234657019376755) & (yn >= -0.39097706973552704), T(-0.2469314038753510), yn)
    yn = torch.where((yn < -0.39097706973552704) & (yn >= -0.7675113677978516), T(-0.5350227355957031), yn)
    yn = torch.where(yn < -0.7675113677978516, T(-1.0), yn)
    yp = (yp - qp) + qp
    yn = (yn - qn) + qn
    return (yp * sp + yn * sn).reshape(shp)


def make_weight(kind, rows, cols, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    if kind == "randn":
        w = torch.randn(rows, cols, generator=g) * 0.02
    elif kind == "uniform":
        w = torch.rand(rows, cols, generator=g) - 0.5
    elif kind == "heavy":
        w = torch.randn(rows, cols, generator=g) ** 3 * 0.02
    else:
        raise ValueError(kind)
    return w.to(dtype)


def tiny_llama(seed=0, dtype=torch.bfloat16):
    """Random-init Llama (2 layers, hidden 256) whose every parameter is drawn from a seeded CPU generator in
    sorted-name order, so the GPU box rebuilds the identical model without shipping a state dict."""
    from transformers import LlamaConfig, LlamaForCausalLM
    cfg = LlamaConfig(hidden_size=256, intermediate_size=512, num_hidden_layers=2, num_attention_heads=4,
                      num_key_value_heads=2, vocab_size=128, max_position_embeddings=64, tie_word_embeddings=False,
                      attn_implementation="eager")
    model = LlamaForCausalLM(cfg)
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in sorted(model.named_parameters()):
            if p.dim() == 1:
                p.copy_(1.0 + 0.1 * torch.randn(p.shape, generator=g))
            else:
                p.copy_(torch.randn(p.shape, generator=g) * (0.05 if "embed" in name or "lm_head" in name else 0.03))
    return model.to(dtype)


QL_GRADS = ["model.layers.0.self_attn.q_proj.weight", "model.layers.1.self_attn.k_proj.weight",
            "model.layers.0.mlp.gate_proj.weight", "model.layers.1.mlp.down_proj.weight"]
