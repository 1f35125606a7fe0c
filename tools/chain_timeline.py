"""Where a phase of the persistent decode chain spends its time: %globaltimer marks written by thread 0 of every CTA
(bdq_gemv_w2_chain_set_prof): 0 phase start, 1 first weight loads issued, 2 dependency wait over, 3 x staged,
4 consumed + reduced + stored, 5 signalled.  Prints medians over CTAs per phase kind and the critical path
(last CTA's signal of phase p-1 -> this CTA's wait over).   python tools/chain_timeline.py [lookahead]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import _lib as L  # noqa: E402
from tools import bench_gemv  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    la = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    ops = bench_gemv.build(dev, batch=1, chain=True)
    ch = bench_gemv.chain_of(ops)
    n = ch.n
    grid = torch.cuda.get_device_properties(dev).multi_processor_count * 2
    buf = torch.zeros(grid * n * 8, dtype=torch.int64, device=dev)
    for _ in range(3):
        ch.run(la)
    torch.cuda.synchronize()
    L.lib().bdq_gemv_w2_chain_set_prof(L.ptr(buf), L.i64(buf.numel()))
    ch.run(la)
    torch.cuda.synchronize()
    L.lib().bdq_gemv_w2_chain_set_prof(None, L.i64(0))
    t = buf.cpu().numpy().reshape(grid, n, 8).astype(np.float64)
    t0 = t[:, 0, 0].min()
    t = t - t0
    print(f"token: {(t[:, n - 1, 5].max()) / 1e3:.1f} us for {n} phases, lookahead {la}")
    names = ["q/k/v", "o", "gate/up", "down"]
    for k in range(4):
        sel = np.arange(4 + k, n, 4)   # skip layer 0
        seg = lambda a, b: np.median(t[:, sel, b] - t[:, sel, a])
        last_sig_prev = t[:, sel - 1, 5].max(axis=0)            # last CTA's signal of the previous phase
        wait_over = np.median(t[:, sel, 2], axis=0)
        first_start = t[:, sel, 0].min(axis=0)
        phase_len = t[:, sel, 5].max(axis=0) - t[:, sel - 1, 5].max(axis=0)
        print(f"{names[k]:8s} phase {np.median(phase_len):7.0f} ns | start->issued {seg(0, 1):6.0f}  issued->wait over {seg(1, 2):6.0f}  "
              f"x staged {seg(2, 3):6.0f}  consume {seg(3, 4):6.0f}  signal {seg(4, 5):6.0f} | "
              f"last signal(p-1) -> median wait over {np.median(wait_over - last_sig_prev):6.0f}  "
              f"skew of signals (max-median) {np.median(t[:, sel, 5].max(axis=0) - np.median(t[:, sel, 5], axis=0)):6.0f}  "
              f"consume p95 {np.percentile(t[:, sel, 4] - t[:, sel, 3], 95):6.0f}")


if __name__ == "__main__":
    main()
