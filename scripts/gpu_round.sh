#!/bin/bash
# One GPU-box pass: GPU tests, bench line, ncu launch list of the same command, full ncu capture of the
# tensor-core decoder kernels.  Outputs under gpurun_out/ (copy what is to be judged into profiles/).
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/bench.json
timeout 600 python scripts/profile_decoder.py 383888 115844 > gpurun_out/profile_decoder.log 2>&1; echo "profile_decoder rc=$?"; cat gpurun_out/profile_decoder.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_decoder_ -c 4 -o gpurun_out/decoder_tc -f \
  python scripts/profile_decoder.py 383888 115844 > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 400 python scripts/bench_configs.py > gpurun_out/configs.jsonl 2> gpurun_out/configs.err; echo "configs rc=$?"
ls -la gpurun_out
