"""run_clip(sharded=True) under torchrun: every rank must end with the clip list and the clipped model of the
single-process run_clip (SURVEY.md 8e autoclip row + 8f N4).

  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/check_run_clip_sharded.py
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200.autoclip import run_clip  # noqa: E402


def tiny(seed, layers=6):
    from transformers import LlamaConfig, LlamaForCausalLM
    cfg = LlamaConfig(vocab_size=160, hidden_size=256, intermediate_size=768, num_hidden_layers=layers, num_attention_heads=4,
                      num_key_value_heads=4, max_position_embeddings=64, tie_word_embeddings=False)
    torch.manual_seed(seed)
    return LlamaForCausalLM(cfg).bfloat16().eval()


def main():
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    cfg = {"zero_point": True, "q_group_size": 128, "quant_type": "int"}
    samples = torch.randint(0, 160, (16, 64), generator=torch.Generator().manual_seed(5))
    a, b = tiny(3), tiny(3)
    ra = run_clip(a, None, 2, cfg, samples=samples, device=dev, sharded=True)
    rb = run_clip(b, None, 2, cfg, samples=samples, device=dev)
    ok = [n for n, _, _ in ra["clip"]] == [n for n, _, _ in rb["clip"]]
    ok = ok and all(torch.equal(x[1], y[1]) and torch.equal(x[2], y[2]) for x, y in zip(ra["clip"], rb["clip"]))
    ok = ok and all(torch.equal(p, q) for p, q in zip(a.parameters(), b.parameters()))
    t = torch.tensor([int(ok)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(json.dumps({"world": world, "entries": len(ra["clip"]), "sharded_run_clip_equals_single_process_on_every_rank": bool(t.item())}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
