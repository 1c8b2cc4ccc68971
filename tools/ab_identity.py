#!/usr/bin/env python
"""Bit-identity of two builds of libpt_b200.so on a GPU box: every output the tests look at (accumulators and BGRA of
whole renders on worlds 1-5 with and without the thin lens and on a synthetic scene with boxes and quads, primary-hit
records, traced rays and paths), plus frame times.  One process per build (the library is chosen at import):
    python tools/ab_identity.py dump <lib.so> <out.npz>      python tools/ab_identity.py compare a.npz b.npz
Test/measurement infrastructure, not product."""
import json, os, sys
import numpy as np
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def dump(lib, out):
    os.environ["PT_B200_LIB"] = os.path.abspath(lib)
    from pathtracer_b200 import host, render
    import synth
    res = {}
    worlds = {w: host.HostWorld(w) for w in (1, 2, 3, 4, 5)}
    rc = render.RenderCore(0)
    W, H, pp = 192, 108, 5
    for w, lens in [(1, False), (2, False), (3, False), (4, False), (5, False), (3, True), (5, True), (1, True)]:
        sc = worlds[w].scene(W, H, lens=lens)
        rc.upload_scene(sc)
        rc.render(W, H, pp, seed=3)
        b, a, st = rc.readback(True, True)
        tag = f"w{w}{'L' if lens else ''}"
        res[tag + "_acc"], res[tag + "_bgra"], res[tag + "_rays"] = a.copy(), b.copy(), np.array([st.rays, st.nan_retries])
        if not lens:
            res[tag + "_prim"] = rc.primary_hits(160, 90).copy()
            rng = np.random.default_rng(w)
            cam = np.array([sc.camera.pos.x, sc.camera.pos.y, sc.camera.pos.z], np.float32)
            o = (cam + rng.normal(0, 0.3, (4000, 3))).astype(np.float32)
            d = rng.normal(size=(4000, 3)).astype(np.float32)
            d /= np.linalg.norm(d, axis=1, keepdims=True)
            res[tag + "_rays_hit"] = rc.trace_rays(o, d).copy()
            res[tag + "_paths"] = rc.trace_paths(o, d, 5).copy()
    ss = synth.SynthScene(seed=4, width=W, height=H)
    rc.upload_scene(ss.scene)
    rc.render(W, H, pp, seed=3)
    b, a, st = rc.readback(True, True)
    res["synth_acc"], res["synth_bgra"] = a.copy(), b.copy()
    o, d = ss.random_rays(6000)
    res["synth_rays_hit"] = rc.trace_rays(o, d).copy()
    res["synth_paths"] = rc.trace_paths(o, d, 5).copy()
    # timings (device events): headline frame at -p16, and the other worlds
    times = {}
    for w, (W2, H2, pp2) in {5: (1920, 1080, 16), 1: (1280, 720, 16), 2: (1280, 720, 16), 3: (1280, 720, 16), 4: (1920, 1080, 8)}.items():
        rc.upload_scene(worlds[w].scene(W2, H2))
        ms = []
        for f in range(3):
            rc.render(W2, H2, pp2, seed=1 + f)
            ms.append(rc.readback(False, False)[2].render_ms)
        times[f"w{w}"] = dict(ms=ms, gsamples_per_s=W2 * H2 * pp2 * pp2 / min(ms[1:]) / 1e6)
    rc.close()
    np.savez(out, **res)
    json.dump(times, open(out + ".times.json", "w"), indent=1)
    print(json.dumps(dict(lib=lib, times=times)))


def compare(a, b):
    A, B = np.load(a), np.load(b)
    bad = []
    for k in A.files:
        x, y = A[k], B[k]
        same = x.shape == y.shape and x.tobytes() == y.tobytes()
        if not same:
            bad.append(k)
    ta, tb = json.load(open(a + ".times.json")), json.load(open(b + ".times.json"))
    print(json.dumps(dict(identical=not bad, differing=bad, n_arrays=len(A.files),
                          speedup={k: tb[k]["gsamples_per_s"] / ta[k]["gsamples_per_s"] for k in ta},
                          a={k: ta[k]["gsamples_per_s"] for k in ta}, b={k: tb[k]["gsamples_per_s"] for k in tb})))
    return 0 if not bad else 1


if __name__ == "__main__":
    sys.exit(dump(sys.argv[2], sys.argv[3]) if sys.argv[1] == "dump" else compare(sys.argv[2], sys.argv[3]))
