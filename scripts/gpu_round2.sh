#!/bin/bash
# Round-2 GPU-box pass.  usage: bash scripts/gpu_round2.sh [quick|full]   (outputs under gpurun_out/)
set -u
MODE=${1:-full}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt; ls -R oracle/_ref > gpurun_out/ref_files.txt 2>&1
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "no-alloc test alone rc=$?"; tail -3 gpurun_out/pytest_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -q --timeout 900 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -40 gpurun_out/pytest_gpu.log | tee gpurun_out/pytest_gpu_tail.txt
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -5 gpurun_out/bench.err
tail -c 6000 gpurun_out/bench.json
ROWS=$(python -c "import json; r=json.load(open('gpurun_out/bench.json'))['rows_per_iteration']; print(r['first_pass'], r['fwd_jac'])" 2>/dev/null || echo "231460 84672")
echo "rows: $ROWS"
timeout 300 python scripts/profile_decoder.py $ROWS > gpurun_out/profile_decoder.log 2>&1; echo "profile_decoder rc=$?"; tail -2 gpurun_out/profile_decoder.log
timeout 600 python scripts/parity_decompose.py > gpurun_out/parity_decompose.md 2> gpurun_out/parity_decompose.err; echo "decompose rc=$?"; cat gpurun_out/parity_decompose.md
for v in "" "QSP_CONCURRENT=0" "QSP_FUSED=0" "QSP_FUSED=0 QSP_CONCURRENT=0"; do
  env $v timeout 300 python scripts/single_ab.py 2>/dev/null | tee -a gpurun_out/single_ab.jsonl
done
QSP_TIMELINE=1 timeout 200 python scripts/timeline.py 2> gpurun_out/timeline.txt; QSP_TIMELINE=1 timeout 200 python scripts/timeline.py kitti_single 2> gpurun_out/timeline_single.txt
tail -25 gpurun_out/timeline_single.txt
if [ "$MODE" = "full" ]; then
  timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference arm rc=$?"; cat gpurun_out/bench_reference.json
  timeout 900 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/plain_bench.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/ncu_bench.log 2>&1; echo "ncu launches rc=$?"
  QSP_CALIBRATE=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_decoder_ -c 4 -o gpurun_out/decoder_tc -f \
    python scripts/profile_decoder.py $ROWS > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
  # launch list of the single-object call (side stream on: QSP_TIMELINE=0), and --set full over the small GN kernels
  QSP_TIMELINE=0 QSP_CALIBRATE=0 timeout 300 python scripts/timeline.py kitti_single > gpurun_out/plain_single.log 2>&1 &&
  QSP_TIMELINE=0 QSP_CALIBRATE=0 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_kitti_single.csv \
    python scripts/timeline.py kitti_single > gpurun_out/ncu_single.log 2>&1; echo "ncu kitti_single launches rc=$?"
  [ -n "${QSP_ROUND_GN_NCU:-}" ] && QSP_TIMELINE=0 QSP_CALIBRATE=0 timeout 900 ncu --set full --clock-control none -k regex:'^k_(prep|sample_fused|phase_list|cand_count|cand_fill|surface_xform|render|accumulate|reduce_partials|solve)' -s 120 -c 22 -o gpurun_out/gn_kernels -f \
    python scripts/timeline.py > gpurun_out/ncu_gn.log 2>&1; echo "ncu gn kernels rc=$?"
fi
ls -la gpurun_out
