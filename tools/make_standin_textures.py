#!/usr/bin/env python
"""Deterministic stand-ins for the three world-1 textures the reference checkout lacks.

The reference's world 1 loads four 2048x2048 PNGs (src/win32_main.cpp:1720-1723); three of them
(`rusty-metal_albedo.png`, `rusty-metal_metallic.png`, `rusty-metal_normal-ogl.png`) are absent
from /root/reference (`.MISSING_LARGE_BLOBS`), and the unmodified reference crashes without them
(src/win32_main.cpp:1601-1602).  This script writes procedural replacements of the same names so
that the reference oracle AND the CUDA build read identical texels.  Every world-1 result in this
repository is therefore "stand-in textures".  `rusty-metal_roughness.png` is the real asset.

Pure integer-hash value noise (tileable at 2048) -> no dependence on numpy's RNG streams.
Usage: python tools/make_standin_textures.py [out_dir=res]   (skips files that already exist)
"""
import hashlib
import os
import sys

import numpy as np
from PIL import Image

SIZE = 2048


def _hash2(ix, iy, seed):
    h = (ix.astype(np.uint32) * np.uint32(374761393)) ^ (iy.astype(np.uint32) * np.uint32(668265263))
    h = h ^ np.uint32((seed * 2246822519) & 0xFFFFFFFF)
    h = (h ^ (h >> np.uint32(13))) * np.uint32(1274126177)
    h = h ^ (h >> np.uint32(16))
    return (h & np.uint32(0xFFFFFF)).astype(np.float32) / np.float32(0xFFFFFF)


def value_noise(cells, seed):
    """Bilinear value noise on a `cells` x `cells` lattice, periodic over SIZE pixels."""
    t = np.arange(SIZE, dtype=np.float32) * np.float32(cells / SIZE)
    i0 = np.floor(t).astype(np.int64)
    f = t - i0.astype(np.float32)
    f = f * f * (3.0 - 2.0 * f)
    i1 = (i0 + 1) % cells
    i0 = i0 % cells
    gx0, gy0 = np.meshgrid(i0, i0)
    gx1, gy1 = np.meshgrid(i1, i1)
    fx, fy = np.meshgrid(f, f)
    a = _hash2(gx0, gy0, seed)
    b = _hash2(gx1, gy0, seed)
    c = _hash2(gx0, gy1, seed)
    d = _hash2(gx1, gy1, seed)
    return (a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy


def fbm(seed, base=8, octaves=5):
    out = np.zeros((SIZE, SIZE), np.float32)
    amp, tot = 1.0, 0.0
    for o in range(octaves):
        out += amp * value_noise(base << o, seed + 17 * o)
        tot += amp
        amp *= 0.5
    return out / tot


def smoothstep(e0, e1, x):
    t = np.clip((x - e0) / (e1 - e0), 0.0, 1.0)
    return t * t * (3.0 - 2.0 * t)


def make_all():
    rust = fbm(101)
    patch = smoothstep(0.48, 0.58, fbm(202, base=4, octaves=4))   # 1 = bare metal, 0 = rust
    height = 0.7 * fbm(303, base=16, octaves=5) + 0.3 * rust

    dark = np.array([0.23, 0.10, 0.06], np.float32)
    orange = np.array([0.70, 0.36, 0.16], np.float32)
    steel = np.array([0.56, 0.57, 0.58], np.float32)
    rustcol = dark[None, None, :] * (1 - rust[..., None]) + orange[None, None, :] * rust[..., None]
    albedo = rustcol * (1 - patch[..., None]) + steel[None, None, :] * patch[..., None]

    metallic = np.repeat((0.05 + 0.9 * patch)[..., None], 3, axis=2)

    scale = 6.0 * SIZE / 256.0
    dx = (np.roll(height, -1, axis=1) - np.roll(height, 1, axis=1)) * 0.5 * scale
    dy = (np.roll(height, -1, axis=0) - np.roll(height, 1, axis=0)) * 0.5 * scale
    n = np.stack([-dx, -dy, np.ones_like(dx)], axis=2)
    n /= np.linalg.norm(n, axis=2, keepdims=True)
    normal = n * 0.5 + 0.5

    def q(a):
        return np.clip(np.floor(a * 255.0 + 0.5), 0, 255).astype(np.uint8)

    return {
        "rusty-metal_albedo.png": q(albedo),
        "rusty-metal_metallic.png": q(metallic),
        "rusty-metal_normal-ogl.png": q(normal),
    }


def main():
    out_dir = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(__file__), "..", "res")
    os.makedirs(out_dir, exist_ok=True)
    names = ["rusty-metal_albedo.png", "rusty-metal_metallic.png", "rusty-metal_normal-ogl.png"]
    if all(os.path.exists(os.path.join(out_dir, n)) for n in names):
        return
    tex = make_all()
    for name, arr in tex.items():
        path = os.path.join(out_dir, name)
        Image.fromarray(arr, "RGB").save(path, optimize=False, compress_level=6)
        print(f"{name}: pixels sha256 {hashlib.sha256(arr.tobytes()).hexdigest()}  file {os.path.getsize(path)} B")


if __name__ == "__main__":
    main()
