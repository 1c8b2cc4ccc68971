#!/bin/bash
# Round 2, GPU call 3 (one B200): the whole GPU suite, resumable dense walks on the triangle tree (sweep), the five worlds with
# the adopted defaults, the parity report at useful spp, bench.py on the headline and on every BASELINE config (N = 1).
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
( timeout 1200 python -m pytest tests -m gpu -q > $O/r02c_pytest_gpu.txt 2>&1; echo "pytest exit $?" >> $O/r02c_pytest_gpu.txt )
tail -6 $O/r02c_pytest_gpu.txt
: > $O/r02c_resume_w5.jsonl

python -c "
import json
for l in open('$O/r02c_resume_w5.jsonl'):
    d = json.loads(l); print('w5 walk_exit', d['walk_exit'], round(d['gsamples_per_s'], 3))"
tools/sweep_worlds.sh "" > $O/r02c_worlds.jsonl 2>&1; cut -c1-200 $O/r02c_worlds.jsonl
timeout 1500 python tools/parity_report.py --budget ${BUDGET:-40} --k 32 --out $O/r02c_parity > $O/r02c_parity.log 2>&1
tail -22 $O/r02c_parity.log
timeout 300 python bench.py > $O/r02c_bench_w5.json 2> $O/r02c_bench_w5.err; tail -c 400 $O/r02c_bench_w5.json
timeout 900 tools/bench_all_configs.sh $O/r02c_bench_configs.jsonl
