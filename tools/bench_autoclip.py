"""Autoclip micro-benchmark (BASELINE config 4): one Llama-2-7B linear per shape, int2-g128, 512 tokens, 100 candidates."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200.autoclip import auto_2clip_layer  # noqa: E402

dev = torch.device("cuda:0")
qt = sys.argv[1] if len(sys.argv) > 1 else "int"
G = int(sys.argv[2]) if len(sys.argv) > 2 else 128   # 64 = the group size of the reference's train scripts
tot_pairs, tot_t = 0, 0.0
for name, co, ci in [("v/o_proj", 4096, 4096), ("gate/up_proj", 11008, 4096), ("down_proj", 4096, 11008)]:
    g = torch.Generator(device=dev).manual_seed(co + ci)
    w = (torch.randn(co, ci, device=dev, generator=g) * 0.02).bfloat16()
    ch = torch.exp(torch.randn(ci, device=dev, generator=g) * 0.5)
    x = (torch.randn(512, ci, device=dev, generator=g) * ch).bfloat16()
    cfg = {"zero_point": True, "q_group_size": G, "quant_type": qt}
    auto_2clip_layer(w, x, 2, cfg)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    auto_2clip_layer(w, x, 2, cfg)
    e1.record()
    torch.cuda.synchronize()
    t = e0.elapsed_time(e1) * 1e-3
    pairs = co * ci // G
    flop = pairs * 2.0 * 512 * G * 101
    print(f"{qt} {name:14s} [{co}x{ci}] {t*1e3:9.2f} ms  {pairs/t/1e6:7.2f} M pairs/s  {flop/t/1e12:7.1f} TFLOP/s executed (tcgen05 for 16-bit g128, else mma.sync)")
    mult = {"v/o_proj": 2, "gate/up_proj": 2, "down_proj": 1}[name]
    tot_pairs += pairs * mult
    tot_t += t * mult
print(f"{qt} g{G} full Llama-2-7B (32 layers x v,o,gate,up,down) on 1 GPU: {tot_t*32:.2f} s ({tot_pairs*32/1e6:.1f} M pairs)")
