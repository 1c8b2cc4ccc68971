#!/usr/bin/env python
"""First-contact GPU check (run under gpurun): parity of the CUDA core against the C oracle on all five
worlds, then a few timed renders.  Writes gpurun_out/gpu_check.json.  Test infrastructure, not product."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)

from oracle import pt_oracle as O  # noqa: E402
from pathtracer_b200 import host, render  # noqa: E402


def main():
    out = {}
    rc = render.RenderCore(0)
    for w in [1, 2, 3, 4, 5]:
        W, H = 320, 180
        hw = host.HostWorld(w)
        sc = hw.scene(W, H)
        rc.upload_scene(sc)
        res = {}
        # primary hits
        o, d = O.primary_rays(sc, W, H)
        ref = O.intersect(sc, o, d)
        gpu = rc.primary_hits(W, H)
        same = (ref["kind"] == gpu["kind"]) & (ref["index"] == gpu["index"])
        res["primary_id_agree"] = float(same.mean())
        hit = ref["kind"] != 0
        res["primary_t_maxrel"] = float(np.max(np.abs(ref["t"][hit & same] - gpu["t"][hit & same]) / ref["t"][hit & same])) if (hit & same).any() else 0.0
        res["primary_mat_agree"] = float((ref["mat"] == gpu["mat"]).mean())
        res["normal_maxabs"] = float(np.max(np.abs(ref["normal"][same] - gpu["normal"][same])))
        # caller-supplied rays
        gpu2 = rc.trace_rays(o, d)
        res["trace_rays_id_agree"] = float(((ref["kind"] == gpu2["kind"]) & (ref["index"] == gpu2["index"])).mean())
        # paths with identical Philox streams
        a = O.raycast(sc, o, d, O.RNG_PHILOX, 3)
        b = rc.trace_paths(o, d, 3)
        fin = np.isfinite(a).all(1) & np.isfinite(b).all(1)
        rel = np.abs(a - b).max(1) / np.maximum(1e-3, np.abs(a).max(1))
        res["paths_close_frac"] = float((rel[fin] < 1e-3).mean())
        res["paths_mean_ref"] = [float(x) for x in a[fin].mean(0)]
        res["paths_mean_gpu"] = [float(x) for x in b[fin].mean(0)]
        res["paths_nonfinite"] = [int((~np.isfinite(a).all(1)).sum()), int((~np.isfinite(b).all(1)).sum())]
        # full render with identical streams
        pp = 4
        W2, H2 = 96, 54
        sc2 = hw.scene(W2, H2)
        rc.upload_scene(sc2)
        rc.render(W2, H2, pp, seed=9)
        bg, acc, st = rc.readback(True, True)
        orc = O.render(sc2, W2, H2, pp, O.RNG_PHILOX, 9)
        gm = acc[..., :3] / np.maximum(acc[..., 3:4], 1)
        om = orc["sum"] / orc["spp"]
        relr = np.abs(gm - om).max(2) / np.maximum(1e-2, np.abs(om).max(2))
        res["render_close_frac"] = float((relr < 2e-3).mean())
        res["render_count_ok"] = bool((acc[..., 3] == pp * pp).all())
        res["render_ldr_equal_frac"] = float((bg == orc["ldr"]).mean())
        res["render_mean_ref"] = [float(x) for x in om.mean((0, 1))]
        res["render_mean_gpu"] = [float(x) for x in gm.mean((0, 1))]
        res["rays_per_sample"] = st.rays / max(1, st.samples)
        res["nan_retries"] = int(st.nan_retries)
        res["bvh_nodes"] = int(st.bvh_nodes)
        out[f"w{w}"] = res
        print(w, json.dumps(res), flush=True)

    # resolve parity on random colours
    rng = np.random.default_rng(1)
    cols = np.concatenate([rng.random((200000, 3)) ** 3 * 4, rng.random((50000, 3)) * 0.01, np.array([[0, 0, 0], [1, 1, 1], [1e9, -1, np.nan]])]).astype(np.float32)
    a = O.resolve(cols)
    b = rc.resolve_colors(cols)
    out["resolve_equal_frac"] = float((a == b).mean())
    diff = np.abs(((a[:, None] >> np.array([0, 8, 16])) & 255).astype(int) - ((b[:, None] >> np.array([0, 8, 16])) & 255).astype(int))
    out["resolve_max_lsb"] = int(diff.max())
    print("resolve", out["resolve_equal_frac"], out["resolve_max_lsb"], flush=True)

    # timing
    timing = {}
    for w, W, H, pp in [(5, 1920, 1080, 8), (5, 1920, 1080, 32), (3, 1280, 720, 16), (2, 1280, 720, 8), (4, 1920, 1080, 8), (1, 1280, 720, 8)]:
        hw = host.HostWorld(w)
        rc.upload_scene(hw.scene(W, H))
        rc.render(W, H, 2)
        rc.synchronize()
        t0 = time.time()
        rc.render(W, H, pp)
        _, _, st = rc.readback(False, False)
        dt = time.time() - t0
        timing[f"w{w}_{W}x{H}_pp{pp}"] = dict(ms=st.render_ms, wall_ms=dt * 1e3, msamples_s=W * H * pp * pp / st.render_ms / 1e3,
                                                rays_per_sample=st.rays / st.samples)
        print(f"w{w} {W}x{H} pp{pp}: {st.render_ms:.2f} ms  {W*H*pp*pp/st.render_ms/1e3:.1f} Msamples/s rays/sample {st.rays/st.samples:.3f}", flush=True)
    out["timing"] = timing
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "gpu_check.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
