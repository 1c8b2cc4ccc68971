#!/usr/bin/env python
"""Small fixed workload for ncu captures and sweeps: uploads a world and renders `--frames` frames.  The scheduling knobs go
through pt_set_tuning (the library reads no environment); PT_B200_LIB selects another build of the library."""
import argparse, json, os, sys
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
from pathtracer_b200 import _ctypes_api as T
from pathtracer_b200 import host, render

ap = argparse.ArgumentParser()
ap.add_argument("--world", type=int, default=5)
ap.add_argument("--width", type=int, default=1920)
ap.add_argument("--height", type=int, default=1080)
ap.add_argument("--pp", type=int, default=8)
ap.add_argument("--frames", type=int, default=2)
ap.add_argument("--kernel", default="auto")
ap.add_argument("--bvh", default="sah", choices=["sah", "lbvh"])
ap.add_argument("--lens", action="store_true")
ap.add_argument("--lanes", type=int, default=-1)
ap.add_argument("--shade-batch", type=int, default=0)
ap.add_argument("--coop-max", type=int, default=-1)
ap.add_argument("--tri-leaf", type=int, default=0)
ap.add_argument("--sph-leaf", type=int, default=0)
ap.add_argument("--json", action="store_true")
a = ap.parse_args()
k = {"auto": T.KERNEL_AUTO, "mega": T.KERNEL_MEGA, "wavefront": T.KERNEL_WAVEFRONT}[a.kernel]
rc = render.RenderCore(0, bvh=T.BVH_DEVICE_LBVH if a.bvh == 'lbvh' else T.BVH_HOST_SAH, kernel=k)
rc.set_tuning(sample_lanes_log2=a.lanes, shade_batch=a.shade_batch, coop_max=a.coop_max, tri_leaf=a.tri_leaf, sph_leaf=a.sph_leaf)
rc.upload_scene(host.HostWorld(a.world).scene(a.width, a.height, lens=a.lens))
ms = []
for i in range(a.frames):
    rc.render(a.width, a.height, a.pp, seed=1 + i)
    _, _, st = rc.readback(False, False)
    ms.append(st.render_ms)
    if not a.json:
        print(f"frame {i}: {st.render_ms:.3f} ms, {a.width*a.height*a.pp*a.pp/st.render_ms/1e3:.1f} Msamples/s, rays/sample {st.rays/max(1,st.samples):.3f}")
if a.json:
    best = min(ms[1:] or ms)
    d = {k: v for k, v in vars(a).items() if k not in ("json", "frames")}
    d.update(ms=best, gsamples_per_s=a.width * a.height * a.pp * a.pp / best / 1e6, rays_per_sample=st.rays / max(1, st.samples),
             lib=os.path.basename(render.LIB_PATH))
    print(json.dumps(d))
