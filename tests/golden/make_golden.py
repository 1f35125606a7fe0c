"""Generate golden vectors by running the UNMODIFIED reference (imported from /root/reference) on CPU.

Run in the build container only:  python tests/golden/make_golden.py
Writes tests/golden/*.npz.  bf16/fp16 tensors are stored as float32 (exact) plus a dtype tag; the
reference has no golden vectors of its own, so these are the parity pin for oracle/.
"""
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference/quantization"
OUT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REF)
sys.modules.setdefault("awq_inference_engine", types.ModuleType("awq_inference_engine"))

sys.path.insert(0, os.path.dirname(OUT))
import quantizer as rq  # noqa: E402
from _util import QL_GRADS, tiny_llama  # noqa: E402

DT = {"f32": torch.float32, "bf16": torch.bfloat16, "f16": torch.float16}


def f32(t):
    return t.detach().to(torch.float32).numpy()


def make_inputs(kind, rows, cols, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    if kind == "randn":
        w = torch.randn(rows, cols, generator=g) * 0.02
    elif kind == "uniform":
        w = torch.rand(rows, cols, generator=g) - 0.5
    elif kind == "heavy":
        w = torch.randn(rows, cols, generator=g) ** 3 * 0.02
    elif kind == "pos":  # groups that do not straddle zero -> z clamps, mask has zeros
        w = torch.rand(rows, cols, generator=g) * 0.1 + 0.3
        w[rows // 2:] = -w[rows // 2:]
    elif kind == "edge":  # constant groups, all-zero groups, tiny ranges, ties
        w = torch.randn(rows, cols, generator=g) * 0.02
        w[0] = 0.0
        w[1] = 0.0123
        w[2] = w[2] * 1e-5
        w[3] = torch.round(w[3] * 200) / 200  # many ties
        w[4, :] = w[4, :].abs()  # min >= 0
        w[5, :] = -w[5, :].abs()
    else:
        raise ValueError(kind)
    return w.to(dtype)


def gen_pq_int():
    out = {}
    idx = 0
    for dname, dtype in DT.items():
        for kind in ["randn", "uniform", "heavy", "pos", "edge"]:
            for n_bit, G in [(2, 128), (2, 64), (3, 128), (4, 128), (8, 128), (2, -1), (4, 32)]:
                w = make_inputs(kind, 8, 256, dtype, 100 + idx)
                dq, s, z = rq.pseudo_quantize_tensor(w.clone(), n_bit=n_bit, zero_point=True, q_group_size=G,
                                                     get_scale_zp=True)
                dq2 = rq.pseudo_quantize_tensor(w.clone(), n_bit=n_bit, zero_point=True, q_group_size=G, inplace=True)
                assert torch.equal(dq, dq2.reshape(dq.shape))
                k = f"c{idx}"
                out[k + "_meta"] = np.array([dname, kind, str(n_bit), str(G)])
                out[k + "_w"], out[k + "_dq"], out[k + "_s"], out[k + "_z"] = f32(w), f32(dq), f32(s), f32(z)
                idx += 1
    np.savez_compressed(os.path.join(OUT, "pq_int.npz"), **out)
    print("pq_int", idx)


def gen_ste():
    out = {}
    idx = 0
    for qname in ["int2", "int3", "nf3"]:
        for dname, dtype in DT.items():
            for kind in ["randn", "heavy", "uniform", "ties", "pos"]:
                for G in [64, 128]:
                    if kind == "ties":
                        w = make_inputs("randn", 8, 256, torch.float32, 300 + idx)
                        lim = w.reshape(8, -1, G).abs().amax(-1, keepdim=True) * 0.7  # emulate apply_clip ties
                        w = torch.clamp(w.reshape(8, -1, G), -lim, lim).reshape(8, 256).to(dtype)
                    else:
                        w = make_inputs(kind, 8, 256, dtype, 300 + idx)
                    if qname == "nf3" and kind == "pos":
                        pass  # all-positive / all-negative groups are fine for nf3
                    gy = torch.randn(8, 256, generator=torch.Generator().manual_seed(900 + idx)).to(dtype)
                    mk = {"int2": lambda: rq.SteInt2AsymQuantizer(G), "int3": lambda: rq.SteInt3AsymQuantizer(G),
                          "nf3": lambda: rq.SteN2F3Quantizer(G)}[qname]
                    res = {}
                    for pname, pd in [("T", dtype), ("f64", torch.float64)]:
                        x = w.to(pd).clone().requires_grad_(True)
                        y = mk()(x)
                        y.backward(gy.to(pd))
                        res[pname] = (y.detach(), x.grad.detach())
                    k = f"c{idx}"
                    out[k + "_meta"] = np.array([qname, dname, kind, str(G)])
                    out[k + "_w"], out[k + "_gy"] = f32(w), f32(gy)
                    out[k + "_y"], out[k + "_gw"] = f32(res["T"][0]), f32(res["T"][1])
                    out[k + "_gw64"] = res["f64"][1].numpy()
                    out[k + "_y64"] = res["f64"][0].numpy()  # where y64 == y the T run and the f64 run took the same codes
                    idx += 1
    # NaN behaviour of nf3 (min exactly 0 in a group)
    w = make_inputs("randn", 2, 256, torch.bfloat16, 77)
    w[0, :128] = w[0, :128].abs()
    w[0, 5] = 0.0
    y = rq.pseudo_quantize_n2f3_tensor(w, 128)
    out["nan_w"], out["nan_y"] = f32(w), f32(y)
    np.savez_compressed(os.path.join(OUT, "ste.npz"), **out)
    print("ste", idx)


def gen_clip_autoclip():
    import autoclip as ra  # pulls transformers/datasets (slow import)
    from clip_utils import apply_clip

    out = {}
    idx = 0
    for dname in ["f32", "bf16"]:
        dtype = DT[dname]
        for qt in ["int", "nf3"]:
            for G, co, ci, ntok in [(128, 64, 256, 512), (64, 64, 128, 1100)]:
                g = torch.Generator().manual_seed(500 + idx)
                w = (torch.randn(co, ci, generator=g) * 0.02).to(dtype)
                ch = torch.exp(torch.randn(ci, generator=g) * 0.5)
                x = (torch.randn(ntok, ci, generator=g) * ch).to(dtype)
                q_config = {"zero_point": True, "q_group_size": G, "quant_type": qt}
                bmax, bmin = ra.auto_2clip_layer(w.clone(), x.clone(), n_bit=2, q_config=q_config)
                k = f"c{idx}"
                out[k + "_meta"] = np.array([dname, qt, str(G)])
                out[k + "_w"], out[k + "_x"] = f32(w), f32(x)
                out[k + "_bmax"], out[k + "_bmin"] = f32(bmax), f32(bmin)
                # apply_clip on a module holding this weight
                lin = torch.nn.Linear(ci, co, bias=False, dtype=dtype)
                lin.weight.data = w.clone()
                holder = torch.nn.Module()
                holder.proj = lin
                torch.nn.Module.cuda = lambda self, *a, **k: self  # the reference calls layer.cuda(); no GPU here
                apply_clip(holder, [("proj", bmax, bmin)])
                out[k + "_wclip"] = f32(holder.proj.weight.data)
                idx += 1
    np.savez_compressed(os.path.join(OUT, "autoclip.npz"), **out)
    print("autoclip", idx)


def gen_pack():
    import qmodule as rm

    out = {}
    idx = 0
    for dname in ["f16", "f32", "bf16"]:
        dtype = DT[dname]
        for w_bit, G, rows, cols in [(2, 128, 32, 256), (2, 128, 16, 384), (2, 64, 16, 256), (4, 128, 16, 256),
                                     (2, 128, 16, 2304), (4, 64, 8, 128)]:
            g = torch.Generator().manual_seed(700 + idx)
            lin = torch.nn.Linear(cols, rows, bias=(idx % 2 == 0), dtype=dtype)
            lin.weight.data = (torch.randn(rows, cols, generator=g) * 0.02).to(dtype)
            w0 = lin.weight.data.clone()
            dq, s, z = rq.pseudo_quantize_tensor(lin.weight.data, n_bit=w_bit, zero_point=True, q_group_size=G,
                                                 get_scale_zp=True)
            lin.weight.data = dq
            ql = rm.WQLinear.from_linear(lin, w_bit, G, False, s, z)
            k = f"c{idx}"
            out[k + "_meta"] = np.array([dname, str(w_bit), str(G)])
            out[k + "_w"], out[k + "_dq"], out[k + "_s"], out[k + "_z"] = f32(w0), f32(dq), f32(s), f32(z)
            out[k + "_qweight"] = ql.qweight.numpy()
            out[k + "_qzeros"] = ql.qzeros.numpy()
            out[k + "_qscales"] = f32(ql.scales)
            idx += 1
    np.savez_compressed(os.path.join(OUT, "pack.npz"), **out)
    print("pack", idx)


def gen_gptq():
    """train/convert_hf_to_gptq_manual.py with its `gptqmodel` import (absent here, used only by a version print) stubbed."""
    import importlib.util
    stub = types.ModuleType("gptqmodel")
    stub.__version__ = "stub"
    stub.GPTQModel = object
    sys.modules.setdefault("gptqmodel", stub)
    spec = importlib.util.spec_from_file_location("ref_gptq", "/root/reference/train/convert_hf_to_gptq_manual.py")
    rg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(rg)

    out = {}
    idx = 0
    for dname in ["bf16", "f16", "f32"]:
        dtype = DT[dname]
        for kind, G, rows, cols in [("randn", 64, 32, 256), ("randn", 128, 48, 640), ("heavy", 128, 32, 1024),
                                    ("edge", 64, 16, 256), ("pos", 32, 32, 128), ("uniform", 16, 16, 64),
                                    ("randn", 256, 16, 512), ("randn", 512, 16, 1024)]:
            lin = torch.nn.Linear(cols, rows, bias=(idx % 2 == 0), dtype=dtype)
            lin.weight.data = make_inputs(kind, rows, cols, dtype, 900 + idx)
            codes, s_t, z_t, bias = rg.get_hf_quant_params(lin, 2, True, G)
            qw, qz, sc, g_idx, b = rg.convert_layer_to_gptq_format(codes, s_t, z_t, bias, 2, G)
            k = f"c{idx}"
            out[k + "_meta"] = np.array([dname, str(G)])
            out[k + "_w"] = f32(lin.weight.data)
            out[k + "_codes"], out[k + "_s_t"], out[k + "_z_t"] = f32(codes), f32(s_t), f32(z_t)
            out[k + "_qweight"], out[k + "_qzeros"] = qw.numpy(), qz.numpy()
            out[k + "_scales"], out[k + "_g_idx"] = f32(sc), g_idx.numpy()
            idx += 1
    np.savez_compressed(os.path.join(OUT, "gptq.npz"), **out)
    print("gptq", idx)


def gen_qlinear():
    """quantization/qlinear.py:5-78 -- the function train/train.py:448 calls: convertModelToQuant on a tiny Llama, one
    forward + loss.backward() on CPU (bf16 compute), logits and four weight gradients stored."""
    import qlinear as rl
    out = {}
    ids = torch.randint(0, 128, (2, 32), generator=torch.Generator().manual_seed(11))
    out["ids"] = ids.numpy()
    for qt in ["int2-asym", "ste-n2f3"]:
        model = tiny_llama()
        model, replaced = rl.convertModelToQuant(model, compute_dtype=torch.bfloat16, quant_type=qt, q_group_size=128)
        assert replaced and isinstance(model.model.layers[0].mlp.up_proj, rl.QLinear)
        assert not isinstance(model.lm_head, rl.QLinear)
        res = model(input_ids=ids, labels=ids)
        res.loss.backward()
        out[qt + "_logits"] = f32(res.logits)
        out[qt + "_loss"] = np.float64(res.loss.item())
        params = dict(model.named_parameters())
        for n in QL_GRADS:
            out[qt + "_g_" + n] = f32(params[n].grad)
    np.savez_compressed(os.path.join(OUT, "qlinear.npz"), **out)
    print("qlinear", {k: v.shape for k, v in out.items() if k.endswith("logits")})


if __name__ == "__main__":
    torch.manual_seed(0)
    which = sys.argv[1:] or ["pq", "ste", "pack", "clip", "gptq", "qlinear"]
    if "pq" in which:
        gen_pq_int()
    if "ste" in which:
        gen_ste()
    if "pack" in which:
        gen_pack()
    if "clip" in which:
        gen_clip_autoclip()
    if "gptq" in which:
        gen_gptq()
    if "qlinear" in which:
        gen_qlinear()
