"""Dev tool: %globaltimer phase timeline of the batch-1 streaming GEMV inside a CUDA graph (BDQ_GEMV_PROF).
usage: python tools/gemv_timeline.py [shape=gate|q|down|layer] [flags]"""
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import _lib as L  # noqa: E402
from tools.bench_gemv import QWEN3_8B, build, run, ctypes_stream  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    which = sys.argv[1] if len(sys.argv) > 1 else "gate"
    flags = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    pre = len(sys.argv) > 3 and sys.argv[3] == "pf"
    xrep = (flags >> 8) & 0xff
    ops = build(dev, layers=36, chain=(which == "layer"))
    if xrep:
        ops = [(o[0].repeat(xrep + 1, 1).contiguous(),) + o[1:] for o in ops]
    fused = False
    if which == "layer":
        sel, fused = ops[: 7 * 12], True
    else:
        name = {"gate": "gate_proj", "q": "q_proj", "down": "down_proj", "k": "k_proj"}[which]
        idx = [i for i, s in enumerate(QWEN3_8B) if s[0] == name][0]
        sel = ops[idx::7][:24]
    lib = L.lib()
    cap = 64 * 1024 * 1024 // 8
    prof = torch.zeros(cap, dtype=torch.int64, device=dev)
    s = torch.cuda.Stream(dev)
    with torch.cuda.stream(s):
        run(sel, flags, ctypes_stream(s), fused, pre)
        torch.cuda.synchronize()
        lib.bdq_gemv_set_prof(ctypes.c_void_p(prof.data_ptr()), ctypes.c_int64(cap))
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=s):
            run(sel, flags | 8, ctypes_stream(s), fused, pre)
        lib.bdq_gemv_set_prof(None, ctypes.c_int64(0))
    for _ in range(3):
        gr.replay()
    torch.cuda.synchronize()
    raw = prof.cpu().numpy().astype(np.uint64)
    p = (raw & np.uint64(0xffffffffff)).astype(np.int64)          # %globaltimer ns (40 bits)
    cyc = (raw >> np.uint64(40)).astype(np.int64)                   # SM clock (24 bits): per-warp deltas only
    nsm = torch.cuda.get_device_properties(dev).multi_processor_count
    grid_of = lambda rb: min(nsm * (1 if flags & 2 else 2), rb)
    # slabs in launch order
    off, launches = 0, []
    i = 0
    while i < len(sel):
        n = 1
        if fused:
            nm = QWEN3_8B[i % 7][0]
            n = 3 if nm == "q_proj" else (2 if nm == "gate_proj" else 1)
        rb = sum(o[7] for o in sel[i:i + n]) // 16
        g = grid_of(rb)
        launches.append((QWEN3_8B[i % 7][0] if fused else which, g, p[off:off + g * 8 * 16].reshape(g, 8, 16),
                         cyc[off:off + g * 8 * 16].reshape(g, 8, 16)))
        off += g * 8 * 16
        i += n
    t0 = min(int(a[:, :, 0].min()) for _, _, a, _ in launches)
    prev_end = None
    print("launch name      grid | first start, median start | median issue-done, wait-done, x-staged | median per-mark "
          "deltas ... | last end  (us, relative to the first launch; gaps vs previous launch's last end)")
    for li, (nm, g, a, cy) in enumerate(launches):
        a = a.astype(np.float64)
        valid = a > 0
        a = np.where(valid, (a - t0) * 1e-3, np.nan)
        st = a[:, :, 0]
        last = np.nanmax(a, axis=2)
        med = lambda v: float(np.nanmedian(v))
        marks = [med(a[:, :, k] - a[:, :, 0]) for k in range(1, 16) if np.isfinite(a[:, :, k]).any()]
        end = float(np.nanmax(last))
        gap = "" if prev_end is None else f" start-prev_end {float(np.nanmin(st)) - prev_end:+6.2f} wait_rel-prev_end {med(a[:, :, 2]) - prev_end:+6.2f}"
        print(f"{li:3d} {nm:10s} {g:4d} | {float(np.nanmin(st)):8.2f} {med(st):8.2f} | dur(max) {end - float(np.nanmin(st)):6.2f} | "
              + " ".join(f"{m:5.2f}" for m in marks) + f" | end {end:8.2f}{gap}")
        prev_end = end
        # fine deltas between consecutive marks from the SM clock (cycles at 1.965 GHz -> ns), median and max over warps
        dc = ((cy[:, :, 1:] - cy[:, :, :-1]) & 0xffffff).astype(np.float64) / 1.965
        ok = valid[:, :, 1:] & valid[:, :, :-1]
        dc = np.where(ok, dc, np.nan)
        nm_ = [k for k in range(dc.shape[2]) if np.isfinite(dc[:, :, k]).any()]
        print("      clk ns med: " + " ".join(f"{np.nanmedian(dc[:, :, k]):5.0f}" for k in nm_) +
              "  | p95: " + " ".join(f"{np.nanpercentile(dc[:, :, k], 95):5.0f}" for k in nm_))


if __name__ == "__main__":
    main()
