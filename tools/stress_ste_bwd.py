"""Randomised stress of the STE backward kernels against the CPU oracle's float64 closed form (the criterion of
tests/test_fakequant_gpu.py:_bwd_check) on adversarial groups: offsets far from zero (the group minimum does not quantise to
code 0 / the maximum not to the top code: the clamp masks of int_thr_bwd_kernel are needed), constant groups, two-valued
groups, tiny and huge ranges, ties at the extremes.   python tools/stress_ste_bwd.py [n_cases]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle  # noqa: E402
from _util import ulp_of  # noqa: E402
from bitdistiller_b200.quantizer import SteInt2AsymQuantizer, SteInt3AsymQuantizer, SteN2F3Quantizer  # noqa: E402

dev = torch.device("cuda:0")
os.environ.setdefault("BDQ_SEMANTICS", "cpu")   # the oracle's closed form is the CPU-eager arithmetic


def adversarial(rows, cols, G, dt, g):
    w = torch.randn(rows, cols, generator=g) * 0.02
    v = w.view(-1, G)
    n = v.shape[0]
    kind = torch.randint(0, 8, (n,), generator=g)
    off = torch.randn(n, 1, generator=g) * 3.0
    v[kind == 1] += off[kind == 1]                                  # far from zero: clamp masks matter
    v[kind == 2] = v[kind == 2][:, :1]                              # constant group
    two = kind == 3
    v[two] = torch.where(torch.rand(int(two.sum()), G, generator=g) > 0.5, v[two][:, :1], v[two][:, 1:2])   # two values
    v[kind == 4] *= 1e-4                                            # tiny range
    v[kind == 5] *= 300.0                                           # large range
    tie = kind == 6
    v[tie, 1] = v[tie].amax(1)                                      # ties at the maximum ...
    v[tie, 2] = v[tie].amin(1)                                      # ... and the minimum
    v[kind == 7] = v[kind == 7].abs() + 0.5                         # all positive, away from zero
    return w.to(dt)


def check(name, mod, oracle_bwd, w, gy, G, dt, fwd_ref=None):
    x = w.to(dev).requires_grad_(True)
    y = mod(x)
    y.backward(gy.to(dev))
    if fwd_ref is not None:    # forward: bit for bit (NaN == NaN, +0 == -0)
        a, b = y.detach().double().cpu(), fwd_ref(w).double()
        nf = int((~((a == b) | (a.isnan() & b.isnan()))).sum())
        if nf:
            print(f"  {name}: {nf} forward values differ from the oracle")
            return nf
    gw = x.grad.double().cpu()
    ref = oracle_bwd(w, gy)
    wg = w.double().reshape(-1, G)
    ext = ((wg == wg.amax(1, keepdim=True)) | (wg == wg.amin(1, keepdim=True))).reshape(ref.shape)
    tol = torch.where(ext, 2.01, 1.01) * ulp_of(ref, dt) + 1e-30
    gsum = gy.double().abs().reshape(-1, G).sum(1, keepdim=True).expand(-1, G).reshape(ref.shape)
    tol = tol + 1e-5 * gsum
    both_nan = gw.isnan() & ref.isnan()
    bad = ((gw - ref).abs() > tol) & ~both_nan
    bad |= gw.isnan() ^ ref.isnan()
    return int(bad.sum())


def main():
    n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 40
    g = torch.Generator().manual_seed(1234)
    total_bad, total = 0, 0
    for case in range(n_cases):
        dt = [torch.bfloat16, torch.float16][case % 2]
        G = [128, 64, 32][case % 3]
        rows, cols = 64 + 32 * (case % 5), [1024, 2048, 512][case % 3]
        w = adversarial(rows, cols, G, dt, g)
        gy = (torch.randn(rows, cols, generator=g)).to(dt)
        for name, mod, ob in (("int2", SteInt2AsymQuantizer(G), lambda a, b: oracle.int_ste_bwd(a, b, G, 2)),
                              ("int3", SteInt3AsymQuantizer(G), lambda a, b: oracle.int_ste_bwd(a, b, G, 3)),
                              ("nf3", SteN2F3Quantizer(G), lambda a, b: oracle.nf3_ste_bwd(a, b, G))):
            fr = {"int2": lambda a: oracle.int_fakequant(a, G, 2, 1)["dq"], "int3": lambda a: oracle.int_fakequant(a, G, 3, 1)["dq"],
                  "nf3": lambda a: oracle.nf3_fakequant(a, G)["dq"]}[name]
            nb = check(name, mod, ob, w, gy, G, dt, fr)
            total_bad += nb
            total += w.numel()
            if nb:
                print(f"case {case} {name} {dt} G={G}: {nb} elements outside the tolerance")
    print(f"{n_cases} cases x 3 quantisers, {total} gradient elements checked (forward values bit for bit as well), {total_bad} outside the tolerance")
    return 1 if total_bad else 0


if __name__ == "__main__":
    sys.exit(main())
