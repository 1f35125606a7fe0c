#!/bin/bash
timeout 300 python tools/gemv_timeline.py layer 1 pf 2>&1 | tail -17
