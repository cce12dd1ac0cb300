#!/bin/bash
# A/B of k_reduce_partials (QSP_REDUCE=0/1): bench step, single-object and pose-only timings
for r in 0 1; do
  echo "QSP_REDUCE=$r"
  QSP_REDUCE=$r python bench.py --no-cpu-baseline | python -c "import json,sys; b=json.loads(sys.stdin.read()); print('  bench', round(b['value'],1), 'objects/s', round(b['ms_per_step'],3), 'ms')"
  QSP_REDUCE=$r python scripts/bench_configs.py 2>/dev/null | head -4 | cut -c1-110
done
