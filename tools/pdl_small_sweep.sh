#!/bin/bash
# which kernel classes should keep programmatic dependent launch in the T <= 8 (draft head) forwards: EB_PDL_SMALL_MASK = class bits
# (common.cuh: GEMM 1, RESID 2, ROPE 4, ATTN 8, COMBINE 16, SWIGLU 32, CONCAT 64, TOPK 128, EXPAND 256)
for m in ${@:-0 1 3 5 9 17 33 65 129 257 511}; do echo "mask=$m: $(EB_PDL_SMALL_MASK=$m python tools/verify_once.py 1 | grep 'round ms' | tail -1)"; done
