#!/bin/bash
# One gpurun call for the warp-cooperative sparse walks: identity + threshold sweep, then the whole GPU test
# suite with the best threshold forced through the environment.  Logs under gpurun_out/.
mkdir -p gpurun_out
rm -f gpurun_out/coop_check.jsonl
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/coop_smi.txt 2>&1
timeout 170 python tools/coop_check.py > gpurun_out/coop_check.log 2>&1
echo "coop_check rc=$?" | tee -a gpurun_out/coop_check.log
BEST=$(python - <<'PY'
import json
best, bt = 0, None
base = None
try:
    for l in open("gpurun_out/coop_check.jsonl"):
        r = json.loads(l)
        if r.get("check") == "timing" and r["world"] == 5:
            t = min(r["ms"][1:] or r["ms"])
            if r["coop_max"] == 0: base = t
            if bt is None or t < bt: bt, best = t, r["coop_max"]
        if r.get("check") == "identical_all" and not r["ok"]:
            best = -1; break
except Exception:
    best = -1
print(best)
PY
)
echo "best coop_max for w5: $BEST" | tee -a gpurun_out/coop_check.log
if [ "$BEST" -gt 0 ]; then
  PT_COOP_MAX=$BEST timeout 330 python -m pytest tests -x -q -m gpu > gpurun_out/coop_pytest_gpu.log 2>&1
  echo "pytest (PT_COOP_MAX=$BEST) rc=$?" | tee -a gpurun_out/coop_pytest_gpu.log
else
  timeout 330 python -m pytest tests -x -q -m gpu > gpurun_out/coop_pytest_gpu.log 2>&1
  echo "pytest (default kernels) rc=$?" | tee -a gpurun_out/coop_pytest_gpu.log
fi
tail -3 gpurun_out/coop_pytest_gpu.log
cat gpurun_out/coop_check.log | tail -40
