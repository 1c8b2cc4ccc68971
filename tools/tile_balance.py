#!/usr/bin/env python
"""Kernel time of every rank's share of one frame (tile sharding), measured on ONE GPU: load balance of the
tile -> rank assignment.  usage: tools/tile_balance.py [G] [world] [W] [H] [pp] [sample_lanes_log2 | -1]"""
import os, sys
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
from pathtracer_b200 import host, render
G = int(sys.argv[1]) if len(sys.argv) > 1 else 8
w = int(sys.argv[2]) if len(sys.argv) > 2 else 5
W = int(sys.argv[3]) if len(sys.argv) > 3 else 1920
H = int(sys.argv[4]) if len(sys.argv) > 4 else 1080
pp = int(sys.argv[5]) if len(sys.argv) > 5 else 32
lanes = int(sys.argv[6]) if len(sys.argv) > 6 else -1
rc = render.RenderCore(0)
rc.set_tuning(sample_lanes_log2=lanes)
rc.upload_scene(host.HostWorld(w).scene(W, H))
rc.render(W, H, pp, seed=1)
_, _, st = rc.readback(False, False)
full = st.render_ms
ms = []
for r in range(G):
    best = 1e9
    for rep in range(2):
        rc.render(W, H, pp, seed=1, tile_rank=r, tile_count=G)
        _, _, st = rc.readback(False, False)
        best = min(best, st.render_ms)
    ms.append(best)
print(f"full frame {full:.2f} ms; ideal share {full / G:.2f} ms")
print("per rank:", " ".join(f"{m:.2f}" for m in ms))
print(f"max {max(ms):.2f} ms = {full / G / max(ms) * 100:.1f} % efficiency; sum {sum(ms):.2f} ms ({sum(ms) / full * 100:.1f} % of the full frame)")
