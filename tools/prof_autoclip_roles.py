"""Who waits for whom inside the tcgen05 autoclip kernel (autoclip_tc.cu): cycles lane 0 of every warp of CTA (0, 0) spent in
its mbarrier waits, out of the role's lifetime.  The role that never waits is the bottleneck."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import _lib as L
from bitdistiller_b200.autoclip import auto_2clip_layer
dev = torch.device("cuda:0")
qt = sys.argv[1] if len(sys.argv) > 1 else "int"
w = (torch.randn(4096, 4096, device=dev) * 0.02).bfloat16()
x = (torch.randn(512, 4096, device=dev) * torch.exp(torch.randn(4096, device=dev) * 0.5)).bfloat16()
cfg = {"zero_point": True, "q_group_size": 128, "quant_type": qt}
auto_2clip_layer(w, x, n_bit=2, q_config=cfg)
buf = torch.zeros(32 * 8, dtype=torch.int64, device=dev)
L.lib().bdq_autoclip_tc_set_prof(L.ptr(buf))
auto_2clip_layer(w, x, n_bit=2, q_config=cfg)
torch.cuda.synchronize()
L.lib().bdq_autoclip_tc_set_prof(None)
t = buf.cpu().reshape(32, 8)
names = {0: ("mma", ["A full", "acc0 empty", "acc1 empty"]), 1: ("param", ["P empty"]), 5: ("rows", ["P full", "A empty"]), 13: ("epi h0", ["acc full"]),
         17: ("epi h1", ["acc full"])}
for wi, (nm, slots) in names.items():
    life = int(t[wi, 7])
    print(f"{nm:8s} (warp {wi:2d}) lifetime {life:9d} cycles = {life/64:.0f} per channel; waits: " +
          ", ".join(f"{s} {int(t[wi, k])/max(life,1)*100:.0f}%" for k, s in enumerate(slots)))
