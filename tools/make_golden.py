#!/usr/bin/env python
"""Generate tests/golden/*.npz from the reference itself (oracle/_ref = the reference's own sources
compiled here).  Runs only in the build container (needs /root/reference); the fixtures travel.

    python tools/make_golden.py            # rewrites tests/golden/

What is pinned, per world 1..5 (scene seed 7 for world 4's layout):
  scene_w*.npz    raw reference scene PODs (materials, spheres, planes, quads, triangles, camera) at 160x90
  hits_w*.npz     RayCastIntersect on the 160x90 jitter-free primary rays: t bits, material, kind, index
  raycast_w*.npz  RayCast radiance (float bits) of those rays with gen.seed(11), sequential draws
  hdr_w*.npz      48x27 render at pp=24 with gen.seed(5): per-pixel sum / sum of squares (double),
                  the reference's own float mean and its packed LDR pixels
plus common.npz: mt19937 uniforms for seed 123, TonemapPass/LinearToSRGB/pack of 4096 colours,
BespokeSampleTexture at 256 points of the four world-1 textures, and a 64x36 test.bmp written by the
reference's WriteDIBImage through its own MasterThread (1 worker, gen.seed(3)).
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
from oracle.ref_harness import RUN_DIR, RefHarness  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
HITS_WH = (160, 90)
HDR_WH, HDR_PP = (48, 27), 24


def main():
    os.makedirs(OUT, exist_ok=True)
    h = RefHarness()
    for w in range(1, 6):
        W, H = HITS_WH
        h.load(w, W, H, 4, 0, 7)
        sc = h.scene()
        np.savez_compressed(os.path.join(OUT, f"scene_w{w}.npz"), **{k: v for k, v in sc.items()})
        o, d = h.primary_rays()
        r = h.intersect(o, d)
        assert (r["kind"] >= 0).all(), "octree != brute force?"
        np.savez_compressed(os.path.join(OUT, f"hits_w{w}.npz"), t=r["t"], mat=r["mat"].astype(np.uint16),
                            kind=r["kind"].astype(np.int8), index=r["index"].astype(np.int32), normal=r["normal"],
                            ray_d_sha=np.frombuffer(hashlib.sha256(d.tobytes()).digest(), np.uint8))
        h.seed(11)
        rgb = h.raycast(o, d)
        np.savez_compressed(os.path.join(OUT, f"raycast_w{w}.npz"), rgb=rgb)
        W2, H2 = HDR_WH
        h.load(w, W2, H2, HDR_PP, 0, 7)
        h.seed(5)
        hd = h.render_hdr()
        np.savez_compressed(os.path.join(OUT, f"hdr_w{w}.npz"), sum=hd["sum"], sumsq=hd["sumsq"], mean_f=hd["mean_f"],
                            ldr=hd["ldr"], pp=np.int32(HDR_PP))
        print("world", w, "done", flush=True)

    # lens (-d) variant on world 3 and the texture-flag variants on world 1
    h.load(3, 48, 27, 8, 1, 7); h.seed(5); hd = h.render_hdr()
    np.savez_compressed(os.path.join(OUT, "hdr_w3_lens.npz"), sum=hd["sum"], sumsq=hd["sumsq"], mean_f=hd["mean_f"], ldr=hd["ldr"], pp=np.int32(8))
    h.load(1, 48, 27, 8, 2 | 4 | 8, 7); h.seed(5); hd = h.render_hdr()
    np.savez_compressed(os.path.join(OUT, "hdr_w1_nmr.npz"), sum=hd["sum"], sumsq=hd["sumsq"], mean_f=hd["mean_f"], ldr=hd["ldr"], pp=np.int32(8))

    h.seed(123)
    uni = h.random(64)
    rng = np.random.default_rng(2024)
    cols = np.concatenate([rng.random((3000, 3)) ** 3 * 6, rng.random((1000, 3)) * 0.02,
                           np.array([[0, 0, 0], [1, 1, 1], [0.0031308, 0.5, 2.0]]), rng.random((93, 3)) * 100]).astype(np.float32)
    res = h.resolve(cols)
    h.load(1, 64, 36, 2, 0, 7)
    uv = ((rng.random((256, 2)) - 0.5) * 40).astype(np.float32)
    tex = np.stack([[h.sample_texture(i, float(u), float(v)) for (u, v) in uv] for i in range(4)])
    # BMP through the reference's own MasterThread + WriteDIBImage
    h.load(3, 64, 36, 2, 0, 7)
    h.seed(3)
    _, px = h.render_threads(1)
    bmp = np.fromfile(os.path.join(RUN_DIR, "test.bmp"), np.uint8)
    np.savez_compressed(os.path.join(OUT, "common.npz"), mt_seed123=uni, resolve_in=cols, resolve_out=res,
                        tex_uv=uv, tex_rgb=tex, bmp_pixels=px, bmp_bytes=bmp)
    total = sum(os.path.getsize(os.path.join(OUT, f)) for f in os.listdir(OUT))
    print("golden bytes:", total)


def make_aabb_fixture():
    """The box primitive (no world holds one): the reference's RayIntersectWithAABB2 over MakeAABB boxes, nearest
    accepted box per ray as RayCastIntersect's loop does (main.cpp:529-553) -> tests/golden/aabb.npz."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from synth import SynthScene
    sc = SynthScene(seed=3, n_boxes=12, n_quads=0, n_spheres=0, plane=False)
    o, d = sc.random_rays(30000, seed=5)
    r = RefHarness().aabb_test(o, d, sc.box_origin, sc.box_half)
    np.savez_compressed(os.path.join(OUT, "aabb.npz"), o=o, d=d, box_origin=sc.box_origin, box_half=sc.box_half,
                        t=r["t"], index=r["index"], face=r["face"])
    print("aabb.npz: hits", int((r["index"] >= 0).sum()), "of", len(o))


if __name__ == "__main__":
    if "--aabb" not in sys.argv:          # --aabb: only (re)write the box fixture
        main()
    make_aabb_fixture()
