"""GPTQ-v2 2-bit export timing: bdq_gptq_export_w2 over all Llama-2-7B projection weights (bf16, G=128).
Algorithmic bytes per weight: 2 read + 0.25 codes + (2 scale + 0.25 zero)/128 written = 2.2676 B."""
import json, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import _lib as L  # noqa: E402
from bitdistiller_b200.convert_hf_to_gptq_manual import convert_layer_to_gptq_format, get_hf_quant_params  # noqa: E402
from tools.bench_gemv import time_fn  # noqa: E402

dev = torch.device("cuda:0")
pk = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "MEASURED_PEAKS.json")
peak = json.load(open(pk)).get("hbm_gbs", 6553.3) if os.path.exists(pk) else 6553.3
G = 128
BPW = 2 + 0.25 + (2 + 0.25) / G
shapes = [(4096, 4096)] * 4 + [(11008, 4096)] * 2 + [(4096, 11008)]
ws = [(torch.randn(oc, ic, device=dev) * 0.02).to(torch.bfloat16) for oc, ic in shapes]
outs = [(torch.empty(ic // 16, oc, dtype=torch.int32, device=dev), torch.empty(ic // G, oc // 16, dtype=torch.int32, device=dev),
         torch.empty(ic // G, oc, dtype=torch.bfloat16, device=dev)) for oc, ic in shapes]
st = L.stream_ptr(dev)


def layer():
    for w, (qw, qz, sc) in zip(ws, outs):
        rc = L.lib().bdq_gptq_export_w2(L.ptr(w), L.ptr(qw), L.ptr(qz), L.ptr(sc), L.i64(w.shape[0]), L.i64(w.shape[1]), G, 2,
                                        L.BF16, st)
        assert rc == 0


nw = sum(a * b for a, b in shapes)
t = time_fn(layer)
print(f"gptq export, one Llama-2-7B layer ({nw/1e6:.1f} M weights, 7 launches, {nw*BPW/1e6:.0f} MB): {t*1e3:.3f} ms  "
      f"{nw*BPW/t/1e9:.0f} GB/s = {nw*BPW/t/1e9/peak*100:.1f}% of HBM peak; 32 layers {t*32*1e3:.1f} ms")
for (oc, ic), w, (qw, qz, sc) in list(zip(shapes, ws, outs))[3:]:
    def one():
        L.lib().bdq_gptq_export_w2(L.ptr(w), L.ptr(qw), L.ptr(qz), L.ptr(sc), L.i64(oc), L.i64(ic), G, 2, L.BF16, st)
    t1 = time_fn(one)
    print(f"  [{oc:5d} x {ic:5d}] {t1*1e6:8.1f} us  {oc*ic*BPW/t1/1e9:7.0f} GB/s (L2-warm for the 4096^2 case)")

# the reference's own code path for the same layer set, on the host cores (imports nothing from /root/reference:
# an eager restatement of its two functions; timing context only)
lin = torch.nn.Linear(4096, 4096, bias=False, dtype=torch.bfloat16)
lin.weight.data = ws[0].cpu()
t0 = time.time()
w32 = lin.weight.data.float().reshape(-1, G)
mx, mn = w32.amax(1, keepdim=True), w32.amin(1, keepdim=True)
s = (mx - mn).clamp(min=1e-5) / 3
z = (-torch.round(mn / s)).clamp_(0, 3)
c = torch.clamp(torch.round(w32 / s) + z, 0, 3).reshape(4096, 4096).to(torch.bfloat16)
ct = c.t().contiguous()
q = torch.zeros((4096 // 16, 4096), dtype=torch.int32)
for i in range(16):
    q |= ct[i::16, :].to(torch.int32) << (i * 2)
t_cpu = time.time() - t0
print(f"eager host restatement of the same arithmetic, one 4096x4096 ({torch.get_num_threads()} threads): {t_cpu*1e3:.0f} ms")
