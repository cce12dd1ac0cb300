#!/bin/bash
# timing experiments with the decoder kernel's debug switches (results are garbage when non-zero)
for d in 0 1 2 3 4 8 9; do
  echo "QSP_TC_DEBUG=$d"; QSP_TC_DEBUG=$d timeout 300 python scripts/profile_decoder.py 383888 115844 2>&1 | tail -1
done
