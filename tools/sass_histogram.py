#!/usr/bin/env python
"""Per-kernel-file SASS opcode evidence for profiles/ (VERDICT r1 item 10): counts of the mnemonics that prove which hardware
path a kernel uses -- tcgen05 (UTCHMMA / UTCBAR / LDTM), TMA (UTMALDG / UBLKCP / UTMAPF), legacy tensor (HMMA / IMMA), shared
vs generic memory (LDS / STS vs LD.E / ST.E), 256-bit global vectors -- from `cuobjdump -sass` of the built objects.
Run in the build container after `make`:  python tools/sass_histogram.py > profiles/r02_sass_opcodes.txt"""
import collections
import glob
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = ["UTCHMMA", "UTCBAR", "LDTM", "UTMALDG", "UBLKCP", "UTMAPF", "SYNCS", "HMMA", "IMMA", "LDS", "STS", "LDG.E.256", "STG.E.256",
        "LD.E", "ST.E", "RED", "ATOM", "LDGSTS", "MEMBAR", "HSET2", "PRMT", "MUFU", "DFMA"]
print(f"{'object':26s} " + " ".join(f"{k:>9s}" for k in KEYS) + "   total")
for obj in sorted(glob.glob(os.path.join(ROOT, "bitdistiller_b200", "csrc", "build", "*.o"))):
    sass = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    ops = collections.Counter()
    tot = 0
    for line in sass.splitlines():
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        tot += 1
        op = m.group(1)
        for k in KEYS:
            if op == k or op.startswith(k + ".") or (k in ("LDG.E.256", "STG.E.256") and op.startswith(k.split(".")[0]) and ".256" in op):
                ops[k] += 1
    print(f"{os.path.basename(obj):26s} " + " ".join(f"{ops[k]:9d}" for k in KEYS) + f"   {tot}")
