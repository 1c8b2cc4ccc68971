#!/usr/bin/env python
"""Wall time of the reference program with the library dropped in (oracle/_ref/pathtracer_ref_b200) against the
reference program alone (oracle/_ref/pathtracer_ref), process start to test.bmp written."""
import os, subprocess, sys, time
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
run = os.path.join(ROOT, "oracle", "_ref", "run")
env = dict(os.environ, PT_REF_SIZE="1920x1080", PT_SEED="1", PT_REF_CORES=str(len(os.sched_getaffinity(0))))
def t(cmd):
    t0 = time.perf_counter()
    r = subprocess.run(cmd, cwd=run, env=env, capture_output=True, text=True)
    return time.perf_counter() - t0, r
for args in (["-w5", "-p32"], ["-w5", "-p32"], ["-w1", "-p16"], ["-w4", "-p23"]):
    s, r = t([os.path.join(run, "..", "pathtracer_ref_b200"), *args])
    print(f"reference program + libpt_b200.so  {' '.join(args)} 1920x1080: {s:.2f} s wall (rc {r.returncode})", flush=True)
n = env["PT_REF_CORES"]
s, r = t([os.path.join(run, "..", "pathtracer_ref"), "-w5", "-p2", f"-t{n}"])
print(f"reference program alone -w5 -p2 -t{n} 1920x1080 (4 spp, 256x fewer samples than -p32): {s:.2f} s wall; {r.stderr.strip().splitlines()[-1] if r.stderr.strip() else ''}")
