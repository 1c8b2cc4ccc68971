#!/usr/bin/env python
"""Times every build/libpt_b200_*.so (tools/build_variants.sh) x sample-lane layouts on one workload.
Each configuration runs in its own process (the library and PT_SAMPLE_LANES_LOG2 are read at load)."""
import argparse, glob, json, os, re, subprocess, sys
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
ap = argparse.ArgumentParser()
ap.add_argument("--world", type=int, default=5)
ap.add_argument("--pp", type=int, default=16)
ap.add_argument("--lanes", default="-1")
ap.add_argument("--libs", default="build/libpt_b200_*.so")
ap.add_argument("--size", default="1920x1080")
ap.add_argument("--out", default="")
ap.add_argument("--kernel", default="")
ap.add_argument("--wfpix", default="-1")
a = ap.parse_args()
W, H = map(int, a.size.split("x"))
rows = []
for lib in sorted(glob.glob(os.path.join(ROOT, a.libs))) or [""]:
    for lanes, wfpix in [(l, x) for l in a.lanes.split(",") for x in a.wfpix.split(",")]:
        env = dict(os.environ)
        if a.kernel: env["PT_KERNEL"] = a.kernel
        if wfpix != "-1": env["PT_WF_PIXELS_LOG2"] = wfpix
        if lib: env["PT_B200_LIB"] = lib
        if lanes != "-1": env["PT_SAMPLE_LANES_LOG2"] = lanes
        r = subprocess.run([sys.executable, os.path.join(ROOT, "tools/profile_render.py"), "--world", str(a.world), "--pp", str(a.pp),
                            "--width", str(W), "--height", str(H), "--frames", "3"], env=env, capture_output=True, text=True, timeout=600)
        ms = [float(m) for m in re.findall(r"frame \d+: ([0-9.]+) ms", r.stdout)]
        row = dict(lib=os.path.basename(lib), kernel=a.kernel, wfpix=int(wfpix), lanes=int(lanes), world=a.world, pp=a.pp, ms=min(ms) if ms else None,
                   gsamples=(W * H * a.pp * a.pp / min(ms) / 1e6) if ms else None, err=r.stderr[-300:] if r.returncode else "")
        rows.append(row)
        print(json.dumps(row), flush=True)
if a.out:
    json.dump(rows, open(a.out, "w"), indent=1)
