#!/bin/bash
# bench.py at N GPUs of one box, back to back (what the driver does for SCALE_rNN.json); $1 = output jsonl, NS = list of N
out=${1:-gpurun_out/scale.jsonl}; shift
: > $out
for n in ${NS:-1 2 4 8}; do
  if [ $n -eq 1 ]; then python bench.py --gpus 1 --steps ${STEPS:-20} --warmup 3 --no-cpu-baseline "$@" | tail -1 >> $out
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) bench.py --gpus $n --steps ${STEPS:-20} --warmup 3 "$@" 2>/dev/null | tail -1 >> $out; fi
done
python - "$out" <<'PY'
import json, sys
rows = [json.loads(l) for l in open(sys.argv[1]) if l.startswith("{")]
for d in rows:
    ph = d.get("phases", {})
    print(f'N={d["n_gpus"]}  {d["value"]/1e3:8.2f} Gsamples/s  e2e {d["e2e"]["value"]/1e3:8.2f}  ms/step {d["ms_per_step"]:8.3f}  shard {d["config"]["shard"]}  '
          f'render ms/rank {ph.get("render_ms_per_rank")}  exchange ms/rank {ph.get("exchange_ms_per_rank")}  clocks {d["clocks"].get("sm_mhz")} {d["clocks"].get("reasons")}')
PY
