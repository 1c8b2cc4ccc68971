#!/bin/bash
# BASELINE.json configs C1..C5 through bench.py, one JSON line each, into $1 (default gpurun_out/configs.jsonl)
out=${1:-gpurun_out/bench_configs.jsonl}
: > $out
python bench.py --world 1 --pp 16 --width 1280 --height 720  --steps 5 --warmup 3 | tail -1 >> $out
python bench.py --world 2 --pp 16 --width 1280 --height 720  --steps 5 --warmup 3 | tail -1 >> $out
python bench.py --world 3 --pp 32 --width 1280 --height 720  --steps 5 --warmup 3 | tail -1 >> $out
python bench.py --world 4 --pp 23 --width 1920 --height 1080 --steps 5 --warmup 3 | tail -1 >> $out
python bench.py --world 5 --pp 32 --width 3840 --height 2160 --steps 3 --warmup 3 | tail -1 >> $out
python - "$out" <<'PY'
import json, sys
for l in open(sys.argv[1]):
    try: d = json.loads(l)
    except Exception: print("bad line:", l[:200]); continue
    cb = d.get("cpu_baseline", {})
    print(f'{d["config"]["workload"]:55s} {d["value"]/1e3:7.2f} Gs/s  e2e {d["e2e"]["value"]/1e3:7.2f}  frac {d["roofline"]["frac"]:.3f}  rays/s {d["roofline"]["rays_per_sample"]:.3f}  cpu {cb.get("value")} ({cb.get("cores")} cores, {cb.get("kind")})')
PY
