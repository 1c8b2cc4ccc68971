#!/usr/bin/env python
"""Create the run directory the reference executable expects.

The reference opens "res\\mario.glb" and "res\\rusty-metal_*.png" relative to its working directory
(src/win32_main.cpp:1459, 1720-1723).  On Linux those are files whose NAME contains a backslash.
They are symlinks to this repository's res/ (real mario.glb + real roughness map + the three
stand-in textures made by tools/make_standin_textures.py).
Usage: python tools/make_run_dir.py <run_dir>
"""
import os
import subprocess
import sys

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
NAMES = ["mario.glb", "rusty-metal_albedo.png", "rusty-metal_metallic.png",
         "rusty-metal_roughness.png", "rusty-metal_normal-ogl.png"]


def main():
    run_dir = os.path.abspath(sys.argv[1])
    os.makedirs(run_dir, exist_ok=True)
    subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "make_standin_textures.py"),
                           os.path.join(ROOT, "res")])
    for n in NAMES:
        src = os.path.join(ROOT, "res", n)
        dst = os.path.join(run_dir, "res\\" + n)
        if os.path.lexists(dst):
            os.remove(dst)
        os.symlink(os.path.relpath(src, run_dir), dst)


if __name__ == "__main__":
    main()
