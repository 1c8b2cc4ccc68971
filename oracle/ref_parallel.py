"""The reference's own render of a whole frame with per-pixel statistics, on every host core.

TEST INFRASTRUCTURE ONLY (tests/, tools/parity_report.py): never imported by the product.

ref_harness.RefHarness.render_hdr is single-threaded (the reference's sampling loops of RenderTexel,
src/win32_main.cpp:1032-1170, around its own RayCast, with per-pixel sum and sum of squares), and the
reference keeps its scene and its std::mt19937 in file-static globals -- so the frame is split into
bands of rows and the bands are rendered by independent worker PROCESSES (`python -m oracle.ref_parallel`),
each with its own copy of the library and its own seed per band.  Where oracle/_ref was never built the
same job runs on the C port (oracle/pt_oracle.c, bit-exact with the reference, mt19937 mode).
"""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
BAND = 4          # rows per band; bands are dealt round-robin so that expensive regions spread over the workers


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def _bands(H, worker, workers):
    return [(y, min(BAND, H - y)) for k, y in enumerate(range(0, H, BAND)) if k % workers == worker]


def _worker(args):
    world, W, H, pp, flags, scene_seed, rng_seed, worker, workers, out = args
    sys.path.insert(0, ROOT)
    from oracle import ref_harness
    rows = _bands(H, worker, workers)
    ys = np.concatenate([np.arange(y0, y0 + hh) for y0, hh in rows]) if rows else np.zeros(0, np.int64)
    at = {int(y): k for k, y in enumerate(ys)}                       # frame row -> row of this worker's compact arrays
    s = np.zeros((len(ys), W, 3), np.float64)
    s2 = np.zeros((len(ys), W, 3), np.float64)
    ldr = np.zeros((len(ys), W), np.uint32)
    if ref_harness.available():
        h = ref_harness.RefHarness().load(world, W, H, pp, flags, scene_seed)
        for y0, hh in rows:
            h.seed((rng_seed * 1000003 + y0) & 0xFFFFFFFF)
            r = h.render_hdr(0, y0, W, hh)
            k = at[y0]
            s[k:k + hh], s2[k:k + hh], ldr[k:k + hh] = r["sum"], r["sumsq"], r["ldr"]
        kind = "reference"
    else:
        from oracle import pt_oracle as O
        from pathtracer_b200 import _ctypes_api as T
        from pathtracer_b200 import host
        tex = T.SCENE_ALL_MAPS & ~((flags >> 1) & 7)
        sc = host.HostWorld(world, scene_seed=scene_seed).scene(W, H, lens=bool(flags & 1), tex_flags=tex)
        for y0, hh in rows:
            r = O.render(sc, W, H, pp, O.RNG_MT19937, (rng_seed * 1000003 + y0) & 0xFFFFFFFF, window=(0, y0, W, hh))
            k = at[y0]
            s[k:k + hh], s2[k:k + hh], ldr[k:k + hh] = r["sum"], r["sumsq"], r["ldr"]
        kind = "port"
    np.savez(out, rows=ys, sum=s, sumsq=s2, ldr=ldr, kind=kind)


def render(world, W, H, pp, flags=0, scene_seed=7, rng_seed=1, workers=None):
    """dict(sum [H,W,3] f64, sumsq, ldr [H,W] u32 BGRA, spp, kind 'reference'|'port', seconds)."""
    import time
    workers = workers or host_cores()
    workers = max(1, min(workers, (H + BAND - 1) // BAND))
    t0 = time.perf_counter()
    with tempfile.TemporaryDirectory() as td:
        procs = []
        for w in range(workers):
            a = [world, W, H, pp, flags, scene_seed, rng_seed, w, workers, os.path.join(td, f"w{w}.npz")]
            procs.append(subprocess.Popen([sys.executable, "-m", "oracle.ref_parallel", json.dumps(a)], cwd=ROOT))
        for p in procs:
            if p.wait() != 0:
                raise RuntimeError("a reference render worker failed")
        s = np.zeros((H, W, 3), np.float64)
        s2 = np.zeros((H, W, 3), np.float64)
        ldr = np.zeros((H, W), np.uint32)
        kind = "reference"
        for w in range(workers):
            z = np.load(os.path.join(td, f"w{w}.npz"))
            ys = z["rows"]
            s[ys], s2[ys], ldr[ys] = z["sum"], z["sumsq"], z["ldr"]
            kind = str(z["kind"])
    return dict(sum=s, sumsq=s2, ldr=ldr, spp=pp * pp, kind=kind, seconds=time.perf_counter() - t0, workers=workers)


def variance_of_mean(r):
    """Per pixel and channel: the squared standard error of the mean, from sum and sum of squares."""
    n = r["spp"]
    m = r["sum"] / n
    return np.maximum(0.0, r["sumsq"] / n - m * m) / max(1, n - 1)


if __name__ == "__main__":
    _worker(json.loads(sys.argv[1]))
