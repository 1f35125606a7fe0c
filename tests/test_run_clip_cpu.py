"""Host logic of run_clip (quantization/autoclip.py:109-207) on CPU: layer-0 input capture, the per-layer hook / forward /
clip order, the skipped q/k projections and the global names -- with the clip-search kernel and apply_clip stubbed (they are
CUDA-only; their parity tests are tests/test_autoclip_gpu.py and tests/test_run_clip_gpu.py)."""
import pytest
import torch


def _tiny_llama(seed):
    from transformers import LlamaConfig, LlamaForCausalLM
    cfg = LlamaConfig(vocab_size=160, hidden_size=256, intermediate_size=768, num_hidden_layers=2, num_attention_heads=4,
                      num_key_value_heads=4, max_position_embeddings=64, tie_word_embeddings=False)
    torch.manual_seed(seed)
    return LlamaForCausalLM(cfg).bfloat16().eval()


def test_run_clip_loop_order_names_and_inputs(monkeypatch):
    import bitdistiller_b200.autoclip as A
    calls, applied = [], []

    def fake_clip(w, x, n_bit, q_config, **kw):
        calls.append((tuple(w.shape), tuple(x.shape), n_bit))
        wg = w.reshape(w.shape[0], -1, q_config["q_group_size"])
        return wg.amax(-1, keepdim=True), wg.amin(-1, keepdim=True)

    monkeypatch.setattr(A, "auto_2clip_layer", fake_clip)
    monkeypatch.setattr(A, "apply_clip", lambda module, cl: applied.append([n for n, _, _ in cl]))
    m = _tiny_llama(0)
    samples = [torch.randint(0, 160, (1, 64)) for _ in range(16)]          # what get_calib_dataset returns
    cfg = {"zero_point": True, "q_group_size": 128, "quant_type": "int"}
    r = A.run_clip(m, None, 2, cfg, samples=samples, device="cpu")
    names = [n for n, _, _ in r["clip"]]
    per_layer = ["self_attn.v_proj", "self_attn.o_proj", "mlp.gate_proj", "mlp.up_proj", "mlp.down_proj"]
    assert names == [f"model.layers.{i}.{n}" for i in range(2) for n in per_layer]
    assert applied == [per_layer, per_layer]
    assert [c[1] for c in calls] == [(16, 64, 256)] * 4 + [(16, 64, 768)] + [(16, 64, 256)] * 4 + [(16, 64, 768)]
    assert all(c[2] == 2 for c in calls)
    assert tuple(r["clip"][4][1].shape) == (256, 6, 1)
    assert not isinstance(m.model.layers[0], torch.nn.Module) or m.model.layers[0].__class__.__name__ == "LlamaDecoderLayer"


def test_get_calib_dataset_local_file_and_hub_names(tmp_path):
    from bitdistiller_b200.clip_utils import append_str_prefix, get_calib_dataset

    class Tok:
        def encode(self, s):
            return [ord(c) % 97 for c in s]

    p = tmp_path / "calib.txt"
    p.write_text("\n".join("sample %d %s" % (i, "x" * (i % 7)) for i in range(40)))
    blocks = get_calib_dataset(str(p), tokenizer=Tok(), n_samples=32, block_size=16)
    assert len(blocks) > 4 and all(tuple(b.shape) == (1, 16) for b in blocks)
    with pytest.raises(NotImplementedError):
        get_calib_dataset("pile", tokenizer=Tok())
    assert append_str_prefix([("a", 1, 2.0)], "m.") == [("m.a", 1, 2.0)]
