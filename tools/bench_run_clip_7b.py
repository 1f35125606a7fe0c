"""End-to-end run_clip (quantization/autoclip.py:109-207 through this package's drop-in) on a random-init Llama-2-7B
architecture with synthetic calibration tokens (no network: no checkpoint, no dataset), int2-g128, on one B200:
layer-0 capture, 32 x (block forward with GPU-resident hooks + clip search of v/o/gate/up/down + apply_clip).

  python tools/bench_run_clip_7b.py [n_samples] [seqlen] [layers]
"""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200.autoclip import run_clip  # noqa: E402


def main():
    n_samples = int(sys.argv[1]) if len(sys.argv) > 1 else 128
    seqlen = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
    layers = int(sys.argv[3]) if len(sys.argv) > 3 else 32
    from transformers import LlamaConfig, LlamaForCausalLM
    dev = torch.device("cuda:0")
    cfg = LlamaConfig(vocab_size=32000, hidden_size=4096, intermediate_size=11008, num_hidden_layers=layers,
                      num_attention_heads=32, num_key_value_heads=32, max_position_embeddings=4096, tie_word_embeddings=False)
    torch.manual_seed(0)
    t0 = time.time()
    with torch.device(dev):
        model = LlamaForCausalLM(cfg).to(torch.bfloat16).eval()
    t_build = time.time() - t0
    samples = torch.randint(0, 32000, (n_samples, seqlen), generator=torch.Generator().manual_seed(1))
    q_config = {"zero_point": True, "q_group_size": 128, "quant_type": "int"}
    torch.cuda.synchronize()
    t0 = time.time()
    res = run_clip(model, None, 2, q_config, n_samples=n_samples, seqlen=seqlen, samples=samples, device=dev)
    torch.cuda.synchronize()
    t = time.time() - t0
    pairs = sum(mx.numel() for _, mx, _ in res["clip"])
    shrunk = sum(int((mx < 0).sum()) for _, mx, _ in res["clip"])
    print(json.dumps({"what": "run_clip, random-init Llama-2-7B architecture, int2-g128, synthetic tokens, 1 GPU, wall clock incl. "
                              "the block forwards", "layers": layers, "tokens": n_samples * seqlen, "entries": len(res["clip"]),
                      "pairs": pairs, "seconds": t, "model_build_seconds": t_build,
                      "peak_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30, "negative_max_vals": shrunk}), flush=True)


if __name__ == "__main__":
    main()
