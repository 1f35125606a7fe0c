#!/bin/bash
for f in 1 1793 16129; do   # flags: pdl | xrep<<8  (xrep 0, 7 -> 8 replicas, 63 -> 64 replicas)
  echo "=== gate flags=$f"; timeout 300 python tools/gemv_timeline.py gate $f 2>&1 | tail -3
  echo "=== q flags=$f"; timeout 300 python tools/gemv_timeline.py q $f 2>&1 | tail -3
done
