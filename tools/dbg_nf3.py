import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
os.environ["BDQ_SEMANTICS"] = "cuda"
import torch
from _util import make_weight, torch_nf3_fakequant
from bitdistiller_b200 import quantizer as Q
import oracle
dt = torch.bfloat16
w = make_weight("randn", 256, 2048, dt, 99)
wd = w.cuda()
e = torch_nf3_fakequant(wd, 128).cpu()
dq = Q.nf3_fakequant(wd, 128)[0].cpu()
o0 = oracle.nf3_fakequant(w, 128, sem=0)["dq"]
o1 = oracle.nf3_fakequant(w, 128, sem=1)["dq"]
print("ours vs eager mismatches", (dq != e).sum().item(), "oracle sem0 vs eager", (o0 != e).sum().item(), "oracle sem1 vs eager", (o1 != e).sum().item())
idx = (dq != e).nonzero()[:10]
x = w.reshape(-1, 128)
for r, c in idx.tolist():
    g = c // 128
    grp = w[r, g*128:(g+1)*128].float()
    mx, mn = grp.max().item(), grp.min().item()
    xv = w[r, c].item()
    s = abs(mx) if xv >= 0 else abs(mn)
    q32 = torch.tensor(xv, dtype=torch.float32) / torch.tensor(s, dtype=torch.float32)
    print(r, c, "x", xv, "s", s, "q32", q32.item(), "qbf16", q32.to(dt).item(), "ours", dq[r, c].item(), "eager", e[r, c].item(), "o0", o0[r,c].item())
# isolate compare semantic on GPU
for thr in [0.8114928305149078, 0.5024898052215576, 0.2826657369732857, 0.0916687622666359, -0.1234657019376755, -0.39097706973552704, -0.7675113677978516]:
    tb = torch.tensor(thr, dtype=torch.float32).to(dt)
    # candidates: bf16 neighbours of thr
    cands = torch.stack([tb, torch.nextafter(tb.float(), torch.tensor(10.0)).to(dt), torch.nextafter(tb.float(), torch.tensor(-10.0)).to(dt)])
    import numpy as np
    bits = tb.view(torch.int16).item()
    cands = torch.tensor([bits - 1, bits, bits + 1], dtype=torch.int16).view(dt)
    print(thr, "cands", cands.float().tolist(), "gpu >=", (cands.cuda() >= thr).tolist(), "cpu >=", (cands >= thr).tolist(), "fp32 >=", (cands.float() >= np.float32(thr)).tolist())
