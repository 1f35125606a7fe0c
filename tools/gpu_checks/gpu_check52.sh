#!/bin/bash
timeout 600 python tools/gemv_variants.py 1 33 1 33 2>&1 | tail -5
