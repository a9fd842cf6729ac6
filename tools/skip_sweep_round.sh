#!/bin/bash
# marginal in-graph cost of each kernel class in the verify step AND in the whole round (EB_SKIP: results are garbage, timing only);
# round delta - verify delta = what the class costs in the ~10 live draft levels
for k in none resid rope combine swiglu attn gemm; do
  out=$(EB_SKIP=$k timeout 200 python tools/verify_once.py 1 2>&1)
  echo "skip=$k: $(echo "$out" | grep 'verify ms' | tail -1) | $(echo "$out" | grep 'round ms' | tail -1)"
done
