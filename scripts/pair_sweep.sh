#!/bin/bash
# CTA-pair (cta_group::2) vs single-CTA decoder kernels: timing, then numerics with pairs on
for p in 0 1; do
  echo "QSP_TC_PAIR=$p"; QSP_TC_PAIR=$p timeout 120 python scripts/profile_decoder.py 383888 115844 2>&1 | tail -2
done
QSP_TC_PAIR=1 timeout 200 python scripts/check_tc.py 2>&1 | head -24
