#!/bin/bash
# Round 2, GPU call 9 (one B200): the GPU suite with the tests added last (4K frame, no-drop, pt_set_camera), smoke(), bench.
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
( timeout 1200 python -m pytest tests -m gpu -q > $O/r02i_pytest_gpu.txt 2>&1; echo "pytest exit $?" >> $O/r02i_pytest_gpu.txt )
tail -5 $O/r02i_pytest_gpu.txt
python -c "import __graft_entry__ as g; g.smoke()" > $O/r02i_smoke.txt 2>&1; cat $O/r02i_smoke.txt
timeout 300 python bench.py > $O/r02i_bench_w5.json 2> $O/r02i_bench_w5.err; tail -c 300 $O/r02i_bench_w5.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/r02i_bench_reference_arm.json 2>&1; tail -c 600 $O/r02i_bench_reference_arm.json
