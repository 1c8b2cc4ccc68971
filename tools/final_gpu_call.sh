#!/bin/bash
# Round-end verification on a B200: threshold sweep of the cooperative walk, the GPU test suite, the bench line.
mkdir -p gpurun_out
timeout 60 python tools/coop_check.py --skip-identity --w5-thresholds 0,8,10,12,14,16,20,24 --out gpurun_out/coop_sweep2.jsonl > gpurun_out/coop_sweep2.log 2>&1
echo "sweep rc=$?"
timeout 175 python -m pytest tests -x -q -m gpu > gpurun_out/final_pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/final_pytest_gpu.log
timeout 120 python bench.py --steps 5 --warmup 3 > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err
echo "bench rc=$?"; cat gpurun_out/final_bench.json; tail -3 gpurun_out/final_bench.err
cat gpurun_out/coop_sweep2.log | tail -12
