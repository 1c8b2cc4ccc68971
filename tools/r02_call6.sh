#!/bin/bash
# Round 2, GPU call 6 (one B200): the whole GPU suite on the final build, per-rank kernel times of the diagonal tile assignment
# (8 shares on one GPU), the headline bench, the ncu launch list of the bench command and ncu --set full of the five variants.
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
COMMIT=${1:-unknown}
( timeout 1200 python -m pytest tests -m gpu -q > $O/r02f_pytest_gpu.txt 2>&1; echo "pytest exit $?" >> $O/r02f_pytest_gpu.txt )
tail -5 $O/r02f_pytest_gpu.txt
{ echo "== C1 w1 1280x720 -p16, 8 shares"; python tools/tile_balance.py 8 1 1280 720 16
  echo "== C2 w2 1280x720 -p16, 8 shares"; python tools/tile_balance.py 8 2 1280 720 16
  echo "== C4 w4 1920x1080 -p8, 8 shares"; python tools/tile_balance.py 8 4 1920 1080 8
  echo "== C5 w5 3840x2160 -p8, 8 shares"; python tools/tile_balance.py 8 5 3840 2160 8
  echo "== headline w5 1920x1080 -p32, 8 shares"; python tools/tile_balance.py 8 5 1920 1080 32
  echo "== headline w5 1920x1080 -p32, 4 shares"; python tools/tile_balance.py 4 5 1920 1080 32; } > $O/r02f_tile_balance.txt 2>&1
cat $O/r02f_tile_balance.txt
timeout 300 python bench.py > $O/r02f_bench_w5.json 2> $O/r02f_bench_w5.err; tail -c 300 $O/r02f_bench_w5.json
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/plain_bench.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02f_launches_bench_w5_1080p.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_bench.log 2>&1
M=smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__thread_inst_executed.sum
for cfg in "5 1920 1080 8" "1 1280 720 8" "2 1280 720 8" "3 1280 720 8" "4 1920 1080 6"; do
  set -- $cfg
  cmd="python tools/profile_render.py --world $1 --width $2 --height $3 --pp $4 --frames 2"
  $cmd > $O/plain_w$1.log 2>&1 &&
  timeout 400 ncu --set full --metrics $M --clock-control none --import-source on -k regex:k_render -s 1 -c 1 -f -o $O/r02f_k_render_w$1 $cmd > $O/ncu_w$1.log 2>&1
  if [ -e $O/r02f_k_render_w$1.ncu-rep ]; then
    python tools/ncu_summary.py $O/r02f_k_render_w$1.ncu-rep $O/r02f_k_render_w$1_summary.txt --commit $COMMIT --samples $(( $2 * $3 * $4 * $4 )) --json $O/k_render_ncu.json --key w$1
    python - $O/k_render_ncu.json w$1 $(( $2 * $3 )) <<'PY'
import json, sys
p, k, px = sys.argv[1], sys.argv[2], int(sys.argv[3])
d = json.load(open(p)); d[k]["pixels"] = px; json.dump(d, open(p, "w"), indent=1, sort_keys=True)
PY
    rm -f $O/r02f_k_render_w$1.ncu-rep
  fi
done
ls $O
