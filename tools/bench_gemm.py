"""Prefill GEMM micro-benchmark (BASELINE config 5b): Qwen3-8B packed int2-g128, M = 2048 tokens.  Both sides (ours, cuBLAS on
the dequantised weights) are timed as 10 calls captured in one CUDA graph: GPU time without Python launch gaps."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import quantizer as Q  # noqa: E402
from bitdistiller_b200.qmodule import pack_wq  # noqa: E402
from bitdistiller_b200.triton_kernels import gemm_w2  # noqa: E402

import json  # noqa: E402
dev = torch.device("cuda:0")
M = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
_pk = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "MEASURED_PEAKS.json")
PEAK = json.load(open(_pk)).get("bf16_tflops_sustained", 1362.8) if os.path.exists(_pk) else 1362.8
tot_f, tot_t = 0.0, 0.0
res = {}
# "qkv fused" = q, k, v packed buffers concatenated along out_features into one WQLinear (fused_attn.py:312-337)
for name, oc, ic, mult in [("q/o_proj", 4096, 4096, 2), ("k/v_proj", 1024, 4096, 2), ("gate/up_proj", 12288, 4096, 2),
                           ("down_proj", 4096, 12288, 1), ("qkv fused", 6144, 4096, 0)]:
    w = (torch.randn(oc, ic, device=dev) * 0.02).half()
    dq, s, z, _ = Q.int_fakequant(w, 2, 128, 0, want_scale_zp=True)
    qw, qz, qs = pack_wq(dq, s.view(oc, -1), z.view(oc, -1), 2, 128)
    x = torch.randn(M, ic, device=dev).half()
    def graph_time(fn, reps=10):   # GPU time per call: `reps` calls in one CUDA graph, median of 10 replays
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(reps):
                fn()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(11)]
        ev[0].record()
        for i in range(10):
            g.replay()
            ev[i + 1].record()
        torch.cuda.synchronize()
        return sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(10))[5] * 1e-3 / reps
    t = graph_time(lambda: gemm_w2(x, qw, qs, qz, 128))
    f = 2.0 * M * oc * ic
    tc = graph_time(lambda: torch.matmul(x, dq.T))
    print(f"M={M} {name:13s} [{oc}x{ic}] {t*1e6:8.1f} us  {f/t/1e12:7.1f} TFLOP/s  ({f/t/1e12/PEAK*100:5.1f}% of measured sustained bf16 peak)"
          f"   cuBLAS fp16 dense on dequantised W: {tc*1e6:8.1f} us {f/tc/1e12:7.1f} TFLOP/s")
    tot_f += f * mult
    tot_t += t * mult
    res[name] = (t, f)
print(f"Qwen3-8B prefill, all 36 layers' linears, M={M}: {tot_t*36*1e3:.2f} ms  {tot_f*36/(tot_t*36)/1e12:.1f} TFLOP/s "
      f"({tot_f/tot_t/1e12/PEAK*100:.1f}% of the sustained bf16 peak {PEAK:.0f})")
tq = res["qkv fused"][0] + res["q/o_proj"][0] + 2 * res["gate/up_proj"][0] + res["down_proj"][0]
print(f"  with q/k/v as one fused WQLinear (4 GEMMs per layer): {tq*36*1e3:.2f} ms  {tot_f/tq/1e12:.1f} TFLOP/s "
      f"({tot_f/tq/1e12/PEAK*100:.1f}%)")
