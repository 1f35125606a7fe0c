"""run_clip (quantization/autoclip.py:109-207) with GPU-resident activation capture.  Checked against the
reference's flow restated with this package's parity-tested pieces: inputs hooked and staged on the HOST (autoclip.py:
170-173), auto_clip_block moving each linear to the GPU and back (:86-107), apply_clip -- same clip list bit for bit and
the same clipped weights."""
import functools
from collections import defaultdict

import pytest
import torch

pytestmark = pytest.mark.gpu


def _tiny_llama(seed):
    from transformers import LlamaConfig, LlamaForCausalLM
    cfg = LlamaConfig(vocab_size=160, hidden_size=256, intermediate_size=768, num_hidden_layers=2, num_attention_heads=4,
                      num_key_value_heads=4, max_position_embeddings=64, tie_word_embeddings=False)
    torch.manual_seed(seed)
    return LlamaForCausalLM(cfg).bfloat16().eval()


@torch.no_grad()
def _host_staged_run_clip(model, samples, w_bit, q_config, dev):
    from bitdistiller_b200.autoclip import auto_clip_block
    from bitdistiller_b200.clip_utils import append_str_prefix, apply_clip, get_named_linears, get_op_name
    model.to(dev)
    layers = model.model.layers
    got = {}

    class _Stop(Exception):
        pass

    def pre(m, args, kw):
        got["hs"], got["kw"] = args[0], dict(kw)
        raise _Stop

    h0 = layers[0].register_forward_pre_hook(pre, with_kwargs=True)
    try:
        model(samples.to(dev))
    except _Stop:
        pass
    h0.remove()
    hs, kwargs = got["hs"], got["kw"]
    res = {"clip": []}
    for layer in layers:
        feat = defaultdict(list)

        def hook(m, x, y, name):
            feat[name].append(x[0].detach().cpu())
        hd = [lin.register_forward_hook(functools.partial(hook, name=n)) for n, lin in get_named_linears(layer).items()]
        out = layer(hs, **kwargs)
        hs = out[0] if isinstance(out, (tuple, list)) else out
        for h in hd:
            h.remove()
        feat = {k: torch.cat(v, dim=0) for k, v in feat.items()}
        cl = auto_clip_block(layer, w_bit=w_bit, q_config=q_config, input_feat=feat)
        layer.to(dev)
        apply_clip(layer, cl)
        res["clip"] += append_str_prefix(cl, get_op_name(model, layer) + ".")
    return res


@pytest.mark.parametrize("quant_type", ["int", "nf3"])
def test_run_clip_resident_equals_host_staged(quant_type):
    from bitdistiller_b200.autoclip import run_clip
    dev = torch.device("cuda:0")
    cfg = {"zero_point": True, "q_group_size": 128, "quant_type": quant_type}
    samples = torch.randint(0, 160, (16, 64), generator=torch.Generator().manual_seed(5))   # 1024 tokens -> stride 2
    a, b = _tiny_llama(3), _tiny_llama(3)
    ra = run_clip(a, None, 2, cfg, samples=samples, device=dev)
    rb = _host_staged_run_clip(b, samples, 2, cfg, dev)
    assert [n for n, _, _ in ra["clip"]] == [n for n, _, _ in rb["clip"]] and len(ra["clip"]) == 10
    for (n, mx, mn), (_, mx2, mn2) in zip(ra["clip"], rb["clip"]):
        assert mx.is_cuda and mx.dtype == torch.bfloat16 and mx.shape[2] == 1
        assert torch.equal(mx, mx2) and torch.equal(mn, mn2), n
    changed = 0
    for (n, pa), (_, pb) in zip(a.named_parameters(), b.named_parameters()):
        assert torch.equal(pa, pb), n
    ref = _tiny_llama(3).to(dev)
    for (n, pa), (_, pr) in zip(a.named_parameters(), ref.named_parameters()):
        changed += int(not torch.equal(pa, pr))
    assert changed == 10      # v, o, gate, up, down of both layers were clipped; q, k, norms, embeddings untouched
