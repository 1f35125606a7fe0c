"""Tiny ncu driver: one auto_2clip_layer call, 4096 x 4096 bf16, 512 tokens, int2-g128."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200.autoclip import auto_2clip_layer
dev = torch.device("cuda:0")
oc = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
w = (torch.randn(oc, 4096, device=dev) * 0.02).bfloat16()
x = torch.randn(512, 4096, device=dev).bfloat16()
for _ in range(2):
    mx, mn = auto_2clip_layer(w, x, n_bit=2, q_config={"zero_point": True, "q_group_size": 128, "quant_type": "int"})
torch.cuda.synchronize()
print("ok", mx.shape)
