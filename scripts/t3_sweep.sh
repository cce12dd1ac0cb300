#!/bin/bash
# cluster-size sweep of the 3-product decoder kernel (timing) + numerics at the default size
for c in 1 2 4; do
  echo "QSP_TC_CLUSTER=$c"; QSP_TC_VERBOSE=1 QSP_TC_CLUSTER=$c timeout 120 python scripts/profile_decoder.py 383888 115844 tensor tensor 2>&1 | tail -2
done
echo "four-product kernel:"; QSP_TC_SPLIT4=1 timeout 120 python scripts/profile_decoder.py 383888 115844 tensor tensor 2>&1 | tail -1
timeout 200 python scripts/check_tc.py 2>&1 | head -16
