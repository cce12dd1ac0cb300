#!/bin/bash
# One GPU-box pass: GPU tests, bench line, ncu launch list of the same command, full ncu capture of the
# tensor-core decoder kernels.  Outputs under gpurun_out/ (copy what is to be judged into profiles/).
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/bench.json
# decoder kernels alone on the row counts of one GN iteration of the bench step (first pass, fwd+Jacobian)
ROWS=$(python -c "import json; r=json.load(open('gpurun_out/bench.json'))['rows_per_iteration']; print(r['first_pass'], r['fwd_jac'])" 2>/dev/null || echo "262144 86016")
echo "rows: $ROWS"
timeout 600 python scripts/profile_decoder.py $ROWS > gpurun_out/profile_decoder.log 2>&1; echo "profile_decoder rc=$?"; cat gpurun_out/profile_decoder.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_decoder_ -c 4 -o gpurun_out/decoder_tc -f \
  python scripts/profile_decoder.py $ROWS > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
# the Gauss-Newton kernels (b)-(d) of one warm iteration under ncu --set full (DRAM bytes and throughput per launch)
timeout 600 ncu --set full --clock-control none -k 'regex:k_(prep|sample|phase_list|candidates|exclusive_scan|surface_xform|refine_scatter|render|row_factors|accumulate|reduce_partials|solve)' \
  --launch-skip 216 -c 20 -o gpurun_out/gn_kernels -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_gn.log 2>&1; echo "ncu gn rc=$?"
QSP_TIMELINE=1 timeout 200 python scripts/timeline.py 2> gpurun_out/timeline.txt; QSP_TIMELINE=1 timeout 200 python scripts/timeline.py kitti_single 2> gpurun_out/timeline_single.txt
timeout 600 python scripts/bench_configs.py > gpurun_out/configs.jsonl 2> gpurun_out/configs.err; echo "configs rc=$?"
ls -la gpurun_out
