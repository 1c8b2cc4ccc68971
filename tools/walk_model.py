#!/usr/bin/env python
"""CPU model of the dense BVH walks of k_render on the Mario scene (no GPU): per-ray node steps of the real tree for
warps of 32 coherent primary rays (a 2x2-pixel block x 8 jittered samples ~ an 8x4 block of a 4x finer pixel grid),
and what handing the last T stragglers of a warp's walk to the cooperative trace would save.
Model: a while-while walk costs the warp max-over-lanes (node steps x 55 + primitive tests x 40) instructions;
a cooperative ray costs 60 + 55 per visited cluster round.  Numbers guide which experiment to spend GPU time on."""
import ctypes as C, os, subprocess, sys
import numpy as np
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import pt_oracle as O          # test infrastructure: used here for the camera rays only
from pathtracer_b200 import host

so = os.path.join(ROOT, "build", "libcluster_check.so")
os.makedirs(os.path.dirname(so), exist_ok=True)
subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-I" + os.path.join(ROOT, "pathtracer_b200", "csrc"),
                       "-I" + os.path.join(ROOT, "include"), "-o", so, os.path.join(ROOT, "tests", "cpp", "cluster_check.cpp")])
L = C.CDLL(so)
hw = host.HostWorld(5)
W, H = 1920 * 4, 1080 * 4
sc = hw.scene(W, H)
tris = host.scene_arrays(sc)["tri_points"].copy()
p = tris.reshape(-1, 3, 3)
boxes = np.ascontiguousarray(np.concatenate([p.min(1), p.max(1)], 1), np.float32)
o, d = O.primary_rays(sc, W, H)
o = o.reshape(H, W, 3); d = d.reshape(H, W, 3)
# region around Mario: find rays that hit the mesh's bounding box via a coarse pass
lo, hi = boxes[:, :3].min(0), boxes[:, 3:].max(0)
inv = 1.0 / np.where(np.abs(d) < 1e-20, 1e-20, d)
t0 = (lo - o) * inv; t1 = (hi - o) * inv
tn = np.minimum(t0, t1).max(2); tf = np.maximum(t0, t1).min(2)
enter = tn <= tf
ys, xs = np.where(enter)
y0, y1, x0, x1 = ys.min() // 4 * 4, (ys.max() // 4 + 1) * 4, xs.min() // 8 * 8, (xs.max() // 8 + 1) * 8
ro = np.ascontiguousarray(o[y0:y1, x0:x1].reshape(-1, 3), np.float32)
rd = np.ascontiguousarray(d[y0:y1, x0:x1].reshape(-1, 3), np.float32)
n = len(ro)
steps, prims, cl, rounds, hit = (np.zeros(n, np.uint32) for _ in range(5))
vp = lambda a: a.ctypes.data_as(C.c_void_p)
rc = L.walk_profile(vp(boxes), vp(tris), len(boxes), 4, n, vp(ro), vp(rd), vp(steps), vp(prims), vp(cl), vp(rounds), vp(hit))
assert rc == 0
hh, ww = y1 - y0, x1 - x0
def warps(a):      # 8x4 blocks of the fine grid
    return a.reshape(hh // 4, 4, ww // 8, 8).transpose(0, 2, 1, 3).reshape(-1, 32)
S, P, CLv, R, ENT = warps(steps), warps(prims), warps(cl), warps(rounds), warps(enter[y0:y1, x0:x1].reshape(-1))
cost_lane = S * 55.0 + P * 40.0
coop_ray = 60.0 + 55.0 * R
k = ENT.sum(1)
dense = k > 12
print(f"warps over Mario's box: {len(S)}, dense (k > 12): {dense.sum()}, mean k {k[dense].mean():.1f}")
print(f"per entering ray: node steps {steps[enter[y0:y1, x0:x1].reshape(-1)].mean():.1f}, prim tests {prims[enter[y0:y1, x0:x1].reshape(-1)].mean():.1f}, "
      f"clusters {cl[enter[y0:y1, x0:x1].reshape(-1)].mean():.2f}, rounds {rounds[enter[y0:y1, x0:x1].reshape(-1)].mean():.2f}")
cd = cost_lane[dense]
srt = np.sort(cd, 1)[:, ::-1]                 # longest first
base = srt[:, 0].sum()
print(f"dense walks: warp cost = max over lanes; mean lane / max lane = {cd.sum() / 32 / srt[:, 0].sum():.2f} (SIMT efficiency of the walk)")
cr = np.take_along_axis(coop_ray[dense], np.argsort(cd, 1)[:, ::-1], 1)        # coop cost of the same lanes, longest first
for T in (0, 1, 2, 3, 4, 6, 8, 12, 16, 32):
    # walk until <= T lanes remain (the (T+1)-th longest lane finishes), then T cooperative rays (with their best t known: upper bound)
    walk = srt[:, T].sum() if T < 32 else 0.0
    tail = cr[:, :T].sum()
    print(f"  tail threshold {T:2d}: walk {walk / base:.3f} + cooperative {tail / base:.3f} = {(walk + tail) / base:.3f} of the present cost")
