#!/bin/bash
# Round 2, GPU call 1 (one B200): the test suite, the headline bench, A/B of the prepared variants (bit-identity + times),
# megakernel vs wavefront kernel per world, tuning sweeps, and ncu --set full of the five k_render variants that ship.
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
COMMIT=${1:-unknown}
( timeout 900 python -m pytest tests -m gpu -x -q > $O/r02a_pytest_gpu.txt 2>&1; echo "pytest exit $?" >> $O/r02a_pytest_gpu.txt ) 
tail -5 $O/r02a_pytest_gpu.txt
timeout 300 python bench.py > $O/r02a_bench_w5.json 2> $O/r02a_bench_w5.err; tail -c 600 $O/r02a_bench_w5.json
timeout 300 python bench.py --world 1 --width 1280 --height 720 --pp 16 --no-cpu-baseline > $O/r02a_bench_c1.json 2> $O/r02a_bench_c1.err
# --- A/B: bit-identity of every output + frame times per world
timeout 200 python tools/ab_identity.py dump pathtracer_b200/libpt_b200.so $O/ab_prod.npz > $O/r02a_ab_prod.log 2>&1
for v in mb7; do
  timeout 200 python tools/ab_identity.py dump build/libpt_b200_$v.so $O/ab_$v.npz > $O/r02a_ab_$v.log 2>&1
  python tools/ab_identity.py compare $O/ab_prod.npz $O/ab_$v.npz > $O/r02a_ab_compare_$v.json; cat $O/r02a_ab_compare_$v.json
  rm -f $O/ab_$v.npz $O/ab_$v.npz.times.json
done
rm -f $O/ab_prod.npz
# --- resumable sphere-tree walks: threshold sweep
: > $O/r02b_resume_sweep.jsonl
for cfg in "2 1280 720 16" "4 1920 1080 12"; do
  set -- $cfg
  for t in 0 2 4 6 8 10 12 16 20 24; do
    timeout 90 python tools/profile_render.py --world $1 --width $2 --height $3 --pp $4 --frames 3 --walk-exit $t --json | tail -1 >> $O/r02b_resume_sweep.jsonl
  done
  for sb in 8 16 24 32; do timeout 90 python tools/profile_render.py --world $1 --width $2 --height $3 --pp $4 --frames 3 --walk-exit 8 --shade-batch $sb --json | tail -1 >> $O/r02b_resume_sweep.jsonl; done
  for l in 2 3 4 5; do timeout 90 python tools/profile_render.py --world $1 --width $2 --height $3 --pp $4 --frames 3 --walk-exit 8 --lanes $l --json | tail -1 >> $O/r02b_resume_sweep.jsonl; done
done
cut -c1-400 $O/r02b_resume_sweep.jsonl | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: print('bad', l[:100]); continue
    print(d['world'], 'walk_exit', d['walk_exit'], 'sb', d['shade_batch'], 'lanes', d['lanes'], round(d['gsamples_per_s'], 3))
"
# --- megakernel vs wavefront kernel, per world (VERDICT: "keep whichever wins per world")
: > $O/r02a_wavefront_ab.jsonl
for cfg in "1 1280 720 16" "2 1280 720 16" "3 1280 720 16" "4 1920 1080 12" "5 1920 1080 16"; do
  set -- $cfg
  for k in mega wavefront; do
    timeout 120 python tools/profile_render.py --world $1 --width $2 --height $3 --pp $4 --frames 3 --kernel $k --json | tail -1 >> $O/r02a_wavefront_ab.jsonl
  done
  PT_B200_LIB=build/libpt_b200_wf_refill.so timeout 120 python tools/profile_render.py --world $1 --width $2 --height $3 --pp $4 --frames 3 --kernel wavefront --json | tail -1 >> $O/r02a_wavefront_ab.jsonl
done
cat $O/r02a_wavefront_ab.jsonl | cut -c1-260
# --- tuning sweeps (pt_set_tuning)
: > $O/r02a_sweeps.jsonl
for sb in 12 16 20 24 28 32; do for cm in 8 12 16; do
  timeout 60 python tools/profile_render.py --world 5 --pp 16 --frames 3 --shade-batch $sb --coop-max $cm --json | tail -1 >> $O/r02a_sweeps.jsonl
done; done
for leaf in 2 3 4 6 8; do timeout 60 python tools/profile_render.py --world 5 --pp 16 --frames 3 --tri-leaf $leaf --json | tail -1 >> $O/r02a_sweeps.jsonl; done
for w in 2 4; do
  for leaf in 1 2 3 4 8; do timeout 60 python tools/profile_render.py --world $w --width 1280 --height 720 --pp 12 --frames 3 --sph-leaf $leaf --json | tail -1 >> $O/r02a_sweeps.jsonl; done
  for cm in 0 4 8 16 24 32; do timeout 60 python tools/profile_render.py --world $w --width 1280 --height 720 --pp 12 --frames 3 --coop-max $cm --json | tail -1 >> $O/r02a_sweeps.jsonl; done
  for sb in 8 16 24 32; do timeout 60 python tools/profile_render.py --world $w --width 1280 --height 720 --pp 12 --frames 3 --shade-batch $sb --json | tail -1 >> $O/r02a_sweeps.jsonl; done
done
# --- config-size parity report
timeout 1500 python tools/parity_report.py --budget ${BUDGET:-25} --k 32 --out $O/r02b_parity > $O/r02b_parity.log 2>&1
tail -20 $O/r02b_parity.log
# --- ncu --set full of the variants that ship (plain run first, as the recipe demands)
M=smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__thread_inst_executed.sum
for cfg in "5 1920 1080 8" "1 1280 720 8" "2 1280 720 8" "3 1280 720 8" "4 1920 1080 6"; do
  set -- $cfg
  cmd="python tools/profile_render.py --world $1 --width $2 --height $3 --pp $4 --frames 2"
  $cmd > $O/plain_w$1.log 2>&1 &&
  timeout 400 ncu --set full --metrics $M --clock-control none --import-source on -k regex:k_render -s 1 -c 1 -f -o $O/r02a_k_render_w$1 $cmd > $O/ncu_w$1.log 2>&1
  if [ -e $O/r02a_k_render_w$1.ncu-rep ]; then
    python tools/ncu_summary.py $O/r02a_k_render_w$1.ncu-rep $O/r02a_k_render_w$1_summary.txt --commit $COMMIT --samples $(( $2 * $3 * $4 * $4 )) --json $O/k_render_ncu.json --key w$1
    python - $O/k_render_ncu.json w$1 $(( $2 * $3 )) <<'PY'
import json, sys
p, k, px = sys.argv[1], sys.argv[2], int(sys.argv[3])
d = json.load(open(p)); d[k]["pixels"] = px; json.dump(d, open(p, "w"), indent=1, sort_keys=True)
PY
    [ $1 -ne 5 ] && [ $1 -ne 4 ] && rm -f $O/r02a_k_render_w$1.ncu-rep
  fi
done
ls -la $O | head -60
