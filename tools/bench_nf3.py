"""NF3 (BASELINE config 3) fake-quant timing on one Llama-2-7B decoder layer's flat bf16 parameter buffer (202.4 M
weights = bench.py's launch shape): forward with and without the fused clip, STE backward.  CUDA events, 3 rotating
buffers (1.2 GB >> L2), median of 20."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import _lib as L  # noqa: E402
try:
    from microbench import timeit  # noqa: E402
except ImportError:
    from tools.microbench import timeit  # noqa: E402


def run(dev, peak, G=128):
    """G = 64 is the group size of the reference's train scripts (scripts/train/train.sh:52), 128 the BASELINE configs'."""
    N = 4 * 4096 * 4096 + 3 * 11008 * 4096
    nbuf = 3
    ws = [(torch.randn(N, device=dev) * 0.02).to(torch.bfloat16) for _ in range(nbuf)]
    gs = [torch.randn(N, device=dev).to(torch.bfloat16) for _ in range(nbuf)]
    out = torch.empty_like(ws[0])
    ng = N // G
    cmax = [(w.view(ng, G).amax(1) * 0.8).contiguous() for w in ws]
    cmin = [(w.view(ng, G).amin(1) * 0.85).contiguous() for w in ws]
    lib, st, n = L.lib(), L.stream_ptr(dev), L.i64(N)

    def nfwd(i):
        lib.bdq_nf3_fakequant_fwd(L.ptr(ws[i]), L.ptr(out), None, None, None, None, None, n, G, 1, 0, st)

    def nclip(i):
        lib.bdq_nf3_fakequant_fwd(L.ptr(ws[i]), L.ptr(out), None, None, None, L.ptr(cmin[i]), L.ptr(cmax[i]), n, G, 1, 0, st)

    def nbwd(i):
        lib.bdq_nf3_ste_bwd(L.ptr(ws[i]), L.ptr(gs[i]), L.ptr(out), n, G, 1, 0, st)

    def ifwd(i):
        lib.bdq_int_fakequant_fwd(L.ptr(ws[i]), L.ptr(out), None, None, None, None, None, n, G, 2, 1, 1, 0, st)

    def ibwd(i):
        lib.bdq_int_ste_bwd(L.ptr(ws[i]), L.ptr(gs[i]), L.ptr(out), n, G, 2, 1, 0, st)

    res = []
    for name, fn, bpw in [("int2 fwd", ifwd, 4), ("int2 ste bwd", ibwd, 6), ("nf3 fwd", nfwd, 4), ("nf3 fwd + clip (config 3)", nclip, 4 + 4 / G),
                          ("nf3 ste bwd", nbwd, 6)]:
        t = timeit(fn, nbuf)
        gbs = N * bpw / t / 1e9
        res.append({"kernel": name, "us_per_layer": t * 1e6, "gbs": gbs, "frac_of_hbm_peak": gbs / peak,
                    "bytes_per_weight": bpw, "weights": N})
    return res


def main():
    dev = torch.device("cuda:0")
    pk = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "MEASURED_PEAKS.json")
    peak = json.load(open(pk)).get("hbm_gbs", 6553.3) if os.path.exists(pk) else 6553.3
    G = int(sys.argv[1]) if len(sys.argv) > 1 else 128
    for r in run(dev, peak, G):
        print(f"bf16 flat layer g{G} {r['kernel']:28s} {r['us_per_layer']:8.1f} us  {r['gbs']:7.1f} GB/s  "
              f"{r['frac_of_hbm_peak']*100:5.1f}% of HBM peak ({peak:.0f})")


if __name__ == "__main__":
    main()
