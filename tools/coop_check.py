#!/usr/bin/env python
"""Warp-cooperative sparse walks (PT_COOP_MAX, k_render<... | PT_F_COOP>) against the default kernel on a GPU box:
(1) accumulators bit-identical on the BVH worlds, (2) frame times for a sweep of the threshold.
Writes gpurun_out/coop_check.jsonl line by line.  Test/measurement infrastructure, not product."""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
from pathtracer_b200 import host, render

ap = argparse.ArgumentParser()
ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "coop_check.jsonl"))
ap.add_argument("--skip-timing", action="store_true")
ap.add_argument("--frames", type=int, default=3)
ap.add_argument("--skip-identity", action="store_true")
ap.add_argument("--w5-thresholds", default="", help="comma list: time only world 5 at these thresholds")
a = ap.parse_args()
os.makedirs(os.path.dirname(a.out), exist_ok=True)
out = open(a.out, "a")


def emit(**kw):
    line = json.dumps(kw)
    print(line, flush=True)
    out.write(line + "\n"); out.flush(); os.fsync(out.fileno())


def core(coop_max):
    if coop_max: os.environ["PT_COOP_MAX"] = str(coop_max)
    else: os.environ.pop("PT_COOP_MAX", None)
    rc = render.RenderCore(0)                  # the environment is read in pt_create
    os.environ.pop("PT_COOP_MAX", None)
    return rc


worlds = {w: host.HostWorld(w) for w in (5, 2, 4)}
ok_all = True
for w, (W, H, pp) in ({} if a.skip_identity else {5: (320, 180, 6), 2: (256, 144, 5), 4: (256, 144, 4)}).items():
    sc = worlds[w].scene(W, H)
    base = core(0); base.upload_scene(sc); base.render(W, H, pp, seed=3)
    _, acc0, st0 = base.readback(False, True); base.close()
    for T in (1, 3, 6, 32):
        c = core(T); c.upload_scene(sc); c.render(W, H, pp, seed=3)
        _, acc1, st1 = c.readback(False, True); c.close()
        same = bool(np.array_equal(acc0.view(np.uint32), acc1.view(np.uint32)))
        ok_all &= same and st0.rays == st1.rays
        emit(check="identical", world=w, coop_max=T, identical=same, rays=[int(st0.rays), int(st1.rays)],
             differing_pixels=int((acc0.view(np.uint32) != acc1.view(np.uint32)).any(2).sum()),
             max_abs=float(np.nanmax(np.abs(acc0 - acc1))))
if not a.skip_identity:
    emit(check="identical_all", ok=bool(ok_all))

if not a.skip_timing:
    plan = [(5, (1920, 1080, 16), (0, 1, 2, 3, 4, 6, 8, 12, 32)), (2, (1280, 720, 16), (0, 2, 4, 8)), (4, (1920, 1080, 8), (0, 2, 4, 8))]
    if a.w5_thresholds:
        plan = [(5, (1920, 1080, 16), tuple(int(t) for t in a.w5_thresholds.split(",")))]
    for w, (W, H, pp), Ts in plan:
        sc = worlds[w].scene(W, H)
        for T in Ts:
            c = core(T); c.upload_scene(sc)
            ms = []
            for f in range(a.frames):
                c.render(W, H, pp, seed=1 + f)
                _, _, st = c.readback(False, False)
                ms.append(st.render_ms)
            c.close()
            emit(check="timing", world=w, size=[W, H], pp=pp, coop_max=T, ms=ms, gsamples_per_s=W * H * pp * pp / min(ms[1:] or ms) / 1e6)
emit(check="done", ok=bool(ok_all))
sys.exit(0 if ok_all else 1)
