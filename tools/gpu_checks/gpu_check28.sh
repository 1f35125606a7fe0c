#!/bin/bash
timeout 600 python tools/gemv_variants.py 1 3 17 19 2>&1 | tail -6
