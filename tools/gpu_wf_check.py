#!/usr/bin/env python
"""Wavefront-in-SM kernel vs the megakernel: same Philox streams -> same samples; only the order of the
per-pixel sums differs.  Prints max relative difference of the pixel means per world, then timings."""
import os, sys
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
import numpy as np
from pathtracer_b200 import _ctypes_api as T
from pathtracer_b200 import host, render

mega = render.RenderCore(0, kernel=T.KERNEL_MEGA)
wf = render.RenderCore(0, kernel=T.KERNEL_WAVEFRONT)
ok = True
for w, lens in [(1, False), (2, False), (3, False), (4, False), (5, False), (3, True), (5, True)]:
    W, H, pp = 200, 120, 6
    sc = host.HostWorld(w).scene(W, H, lens=lens)
    out = []
    for core in (mega, wf):
        core.upload_scene(sc)
        core.render(W, H, pp, seed=3)
        bgra, acc, st = core.readback(True, True)
        out.append((bgra, acc, st))
    (b0, a0, s0), (b1, a1, s1) = out
    m0, m1 = a0[..., :3] / a0[..., 3:4], a1[..., :3] / a1[..., 3:4]
    rel = np.abs(m0 - m1).max(2) / np.maximum(1e-3, np.abs(m0).max(2))
    same_cnt = bool((a0[..., 3] == a1[..., 3]).all())
    print(f"w{w} lens={int(lens)}: max rel {rel.max():.2e}  frac>1e-4 {(rel > 1e-4).mean():.5f}  bytes equal {(b0 == b1).mean():.5f} "
          f"counts equal {same_cnt}  rays {s0.rays} vs {s1.rays}  nan {s0.nan_retries} vs {s1.nan_retries}", flush=True)
    ok &= same_cnt and (rel > 1e-3).mean() < 1e-3 and s0.rays == s1.rays
print("WF_CHECK", "OK" if ok else "FAIL")
for w, (W, H), pp in [(5, (1920, 1080), 16), (4, (1280, 720), 12), (2, (1280, 720), 16), (1, (1280, 720), 16), (3, (1280, 720), 16)]:
    sc = host.HostWorld(w).scene(W, H)
    for name, core in (("mega", mega), ("wavefront", wf)):
        core.upload_scene(sc)
        ms = []
        for i in range(3):
            core.render(W, H, pp, seed=1 + i)
            _, _, st = core.readback(False, False)
            ms.append(st.render_ms)
        print(f"w{w} {W}x{H} pp{pp} {name}: {min(ms):.2f} ms  {W*H*pp*pp/min(ms)/1e6:.2f} Gsamples/s", flush=True)
