#!/bin/bash
# marginal in-graph cost of each kernel class of the verify step (EB_SKIP: results are garbage, timing only)
for k in none resid rope combine swiglu attn gemm "resid,rope,combine,swiglu" "resid,rope,combine,swiglu,attn"; do
  echo "skip=$k: $(EB_SKIP=$k timeout 200 python tools/verify_once.py 1 2>&1 | grep 'verify ms' | tail -1)"
done
