#!/bin/bash
# same for the T > 8 (verify) forward: EB_PDL_BIG_MASK = classes that keep programmatic dependent launch
for m in ${@:-511 509 495 479 507 503 1 9}; do echo "mask=$m: $(EB_PDL_BIG_MASK=$m EB_PDL_SMALL_MASK=1 python tools/verify_once.py 1 | grep 'verify ms' | tail -1)"; done
