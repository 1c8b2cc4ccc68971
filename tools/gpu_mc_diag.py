#!/usr/bin/env python
"""Diagnostic for the Monte-Carlo parity test: z statistics of GPU vs reference fixture, and the same
statistic with the ORACLE (Philox, same spp as the fixture) standing in for the reference (null check)."""
import os, sys
import numpy as np
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
from oracle import pt_oracle as O
from pathtracer_b200 import host, render

def zstats(mu_gpu, se_gpu2, s, s2, n):
    mu = s / n
    var = np.maximum(s2 / n - mu ** 2, 0) * n / (n - 1)
    se = np.sqrt(var / n + se_gpu2)
    ok = se > 1e-9
    z = (mu_gpu - mu)[ok] / se[ok]
    return dict(frac3=float((np.abs(z) > 3).mean()), frac5=float((np.abs(z) > 5).mean()), med=float(np.median(z)), mean=float(z.mean()), std=float(z.std()),
                gmean_ratio=[float(x) for x in (mu_gpu.mean((0, 1)) / np.maximum(mu.mean((0, 1)), 1e-12))], n_ok=int(ok.sum()))

rc = render.RenderCore(0)
W, H = 48, 27
for name, w, lens, flags in [("hdr_w1.npz", 1, False, 7), ("hdr_w2.npz", 2, False, 7), ("hdr_w3.npz", 3, False, 7), ("hdr_w4.npz", 4, False, 7),
                             ("hdr_w5.npz", 5, False, 7), ("hdr_w3_lens.npz", 3, True, 7), ("hdr_w1_nmr.npz", 1, False, 0)]:
    g = np.load(os.path.join(ROOT, "tests", "golden", name))
    pp = int(g["pp"]); n = pp * pp
    sc = host.HostWorld(w).scene(W, H, lens=lens, tex_flags=flags)
    rc.upload_scene(sc)
    means = []
    for seed in range(8):
        rc.render(W, H, pp * 4, seed=1000 + seed)
        _, acc, _ = rc.readback(False, True)
        means.append(acc[..., :3].astype(np.float64) / acc[..., 3:4])
    means = np.stack(means)
    mu_gpu, se_gpu2 = means.mean(0), means.var(0, ddof=1) / len(means)
    ref = zstats(mu_gpu, se_gpu2, g["sum"], g["sumsq"], n)
    orc = O.render(sc, W, H, pp, O.RNG_PHILOX, 4242)
    null = zstats(mu_gpu, se_gpu2, orc["sum"], orc["sumsq"], n)
    O.seed_mt(99)
    orc2 = O.render(sc, W, H, pp, O.RNG_MT19937)
    null2 = zstats(mu_gpu, se_gpu2, orc2["sum"], orc2["sumsq"], n)
    print(name, "\n  vs reference fixture:", ref, "\n  vs oracle philox    :", null, "\n  vs oracle mt seed 99:", null2, flush=True)
