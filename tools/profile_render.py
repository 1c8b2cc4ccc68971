#!/usr/bin/env python
"""Small fixed workload for ncu captures: uploads a world and renders `--frames` frames."""
import argparse, os, sys
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
a = ap.parse_args()
k = {"auto": T.KERNEL_AUTO, "mega": T.KERNEL_MEGA, "wavefront": T.KERNEL_WAVEFRONT}[a.kernel]
rc = render.RenderCore(0, bvh=T.BVH_DEVICE_LBVH if a.bvh == 'lbvh' else T.BVH_HOST_SAH, kernel=k)
rc.upload_scene(host.HostWorld(a.world).scene(a.width, a.height))
for i in range(a.frames):
    rc.render(a.width, a.height, a.pp, seed=1 + i)
    _, _, st = rc.readback(False, False)
    print(f"frame {i}: {st.render_ms:.3f} ms, {a.width*a.height*a.pp*a.pp/st.render_ms/1e3:.1f} Msamples/s, rays/sample {st.rays/max(1,st.samples):.3f}")
