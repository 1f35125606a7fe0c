"""Shared helpers for the parity tests."""
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
DT = {"f32": torch.float32, "bf16": torch.bfloat16, "f16": torch.float16}


def load_cases(name):
    """Yield (meta list[str], dict suffix->np.ndarray) for every case of tests/golden/<name>.npz."""
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    n = len([k for k in z.files if k.endswith("_meta") and k.startswith("c")])
    for i in range(n):
        pre = f"c{i}_"
        yield [str(s) for s in z[pre + "meta"]], {k[len(pre):]: z[k] for k in z.files if k.startswith(pre)}


def t(a, dtype):
    return torch.from_numpy(np.asarray(a)).to(dtype)


def bits_equal(a, b):
    """Value equality that treats NaN==NaN and +0==-0 (eager ops do not pin the sign of zero)."""
    a, b = a.to(torch.float64), b.to(torch.float64)
    return bool(((a == b) | (a.isnan() & b.isnan())).all())


def ulp_of(x, dtype):
    """Size of one ulp of dtype at |x| (elementwise, float64)."""
    x = x.to(torch.float64).abs()
    mant = {torch.float32: 23, torch.bfloat16: 7, torch.float16: 10}[dtype]
    tiny = {torch.float32: -126, torch.bfloat16: -126, torch.float16: -14}[dtype]
    e = torch.floor(torch.log2(torch.clamp(x, min=2.0 ** tiny)))
    return torch.pow(torch.tensor(2.0, dtype=torch.float64), e - mant)
