"""The two executables: `image_compare` (successor of src/image_compare.c) on the CPU, `pathtracer` (the
reference's command line on the CUDA core) on the GPU."""
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT
from pathtracer_b200 import host

BIN = os.path.join(ROOT, "pathtracer_b200", "bin")
REF_CMP = os.path.join(ROOT, "oracle", "_ref", "image_compare_ref")


@pytest.fixture(scope="module", autouse=True)
def _built():
    subprocess.run(["make", "-C", os.path.join(ROOT, "pathtracer_b200", "host")], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    if not os.path.exists(os.path.join(BIN, "image_compare")):
        pytest.skip("bin/image_compare not built")


def _pair(tmp_path):
    rng = np.random.default_rng(3)
    a = rng.integers(0, 256, (40, 64, 4), dtype=np.uint8)
    b = a.copy()
    b[5:20, 8:40, :3] = rng.integers(0, 256, (15, 32, 3), dtype=np.uint8)
    a[..., 3] = b[..., 3] = 255
    pa, pb = str(tmp_path / "a.bmp"), str(tmp_path / "b.bmp")
    host.write_bmp(pa, a.view(np.uint32)[..., 0])
    host.write_bmp(pb, b.view(np.uint32)[..., 0])
    return pa, pb, a, b


def _legacy(a, b):
    """CompareImages of the reference on BGRA bytes (image_compare.c:79-134): no red, alpha twice."""
    d = (a.astype(np.float64) - b.astype(np.float64)) / 255.0            # byte order in memory: B, G, R, A
    dist = np.sqrt(d[..., 3] ** 2 + d[..., 0] ** 2 + d[..., 1] ** 2 + d[..., 3] ** 2)
    return (1.0 - dist).mean() * 100.0


def test_image_compare_legacy_metric_and_rmse(tmp_path):
    pa, pb, a, b = _pair(tmp_path)
    out = subprocess.run([os.path.join(BIN, "image_compare"), pa, pb], capture_output=True, text=True, check=True).stdout
    sim = float(re.search(r"Percentage Similarity: ([0-9.]+) %", out).group(1))
    rmse = float(re.search(r"RMSE \(0-255, RGB\): ([0-9.]+)", out).group(1))
    assert abs(sim - _legacy(a, b)) < 1e-5
    d = a[..., :3].astype(np.float64) - b[..., :3].astype(np.float64)
    assert abs(rmse - np.sqrt((d ** 2).mean())) < 1e-5
    same = subprocess.run([os.path.join(BIN, "image_compare"), pa, pa], capture_output=True, text=True, check=True).stdout
    assert "Percentage Similarity: 100.000000 %" in same and "max abs diff: 0" in same


def test_image_compare_equals_the_reference_program(tmp_path):
    """Same first line as the reference's own ImageCompare built from src/image_compare.c (oracle/_ref)."""
    if not os.path.exists(REF_CMP):
        pytest.skip("oracle/_ref/image_compare_ref not built (needs /root/reference)")
    pa, pb, _, _ = _pair(tmp_path)
    ours = subprocess.run([os.path.join(BIN, "image_compare"), pa, pb], capture_output=True, text=True).stdout.splitlines()[0]
    ref = subprocess.run([REF_CMP, pa, pb], capture_output=True, text=True).stdout.splitlines()[0]
    assert ours == ref


def test_image_compare_errors(tmp_path):
    pa, _, _, _ = _pair(tmp_path)
    r = subprocess.run([os.path.join(BIN, "image_compare"), pa, str(tmp_path / "missing.bmp")], capture_output=True, text=True)
    assert "is an invalid file path" in r.stdout and r.returncode != 0
    small = str(tmp_path / "s.bmp")
    host.write_bmp(small, np.zeros((40, 32), np.uint32))
    r = subprocess.run([os.path.join(BIN, "image_compare"), pa, small], capture_output=True, text=True)
    assert "Images must have same width" in r.stdout
    assert subprocess.run([os.path.join(BIN, "image_compare")], capture_output=True, text=True).stdout.startswith("usage:")


# ---- the render executable -----------------------------------------------------------------------
def _run_cli(tmp_path, *args):
    exe = os.path.join(BIN, "pathtracer")
    if not os.path.exists(exe):
        pytest.skip("bin/pathtracer not built")
    out = str(tmp_path / "out.bmp")
    r = subprocess.run([exe, *args, "--out", out, "--res", os.path.join(ROOT, "res")], capture_output=True, text=True, cwd=str(tmp_path))
    assert r.returncode == 0, r.stderr + r.stdout
    return host.read_bmp(out), r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("w", [3, 5])
def test_cli_matches_the_library(tmp_path, worlds, core, w):
    """`pathtracer -w<k> -p4` writes the same bytes as pt_render through ctypes with the same seed; progressive
    snapshots, the LBVH build and a 2-context tile gather (both on device 0 is not expressible from the CLI,
    so only --gpus 1 here) do not change them."""
    W, H = 160, 96
    core.upload_scene(worlds(w).scene(W, H))
    core.render(W, H, 4, seed=5)
    want, _, _ = core.readback(True, False)
    got, text = _run_cli(tmp_path, f"-w{w}", "-p4", "--width", str(W), "--height", str(H), "--seed", "5")
    assert "Done. Image written to" in text
    assert np.array_equal(got, want)
    got2, text2 = _run_cli(tmp_path, f"-w{w}", "-p4", "--width", str(W), "--height", str(H), "--seed", "5", "--snapshots", "3", "--lbvh")
    assert "snapshot 3/3: 16 of 16 spp" in text2
    assert (got2 == want).mean() > 0.999          # chunked accumulation: float summation order only


@pytest.mark.gpu
def test_cli_fails_loudly(tmp_path):
    exe = os.path.join(BIN, "pathtracer")
    if not os.path.exists(exe):
        pytest.skip("bin/pathtracer not built")
    r = subprocess.run([exe, "-w5", "-p1", "--res", str(tmp_path / "nowhere")], capture_output=True, text=True, cwd=str(tmp_path))
    assert r.returncode != 0 and "cannot build world 5" in r.stderr


# ---- the drop-in itself: the reference program with MasterThread handed to the library ----------
@pytest.mark.gpu
@pytest.mark.parametrize("w,extra", [(5, []), (1, []), (3, ["-d"]), (4, [])])
def test_reference_program_with_the_library_dropped_in(worlds, core, w, extra):
    """oracle/_ref/pathtracer_ref_b200 = the UNMODIFIED reference program (its ParseArgs, LoadWorld, GLB and texture
    loaders, DefineCamera, WriteDIBImage) where the thread shim substitutes MasterThread by the INTEGRATION.md patch
    (g_world / g_camera / g_textures -> pt_upload_scene -> pt_render -> pt_readback into g_image).  Its test.bmp must
    be byte-identical to what the repo's own host + the same library produce for the same seed."""
    exe = os.path.join(ROOT, "oracle", "_ref", "pathtracer_ref_b200")
    run = os.path.join(ROOT, "oracle", "_ref", "run")
    if not (os.path.exists(exe) and os.path.isdir(run)):
        pytest.skip("oracle/_ref/pathtracer_ref_b200 not built (needs /root/reference and libpt_b200.so)")
    W, H, pp = 160, 96, 4
    env = dict(os.environ, PT_REF_SIZE=f"{W}x{H}", PT_SEED="5", PT_REF_SEED="7")     # PT_REF_SEED: world 4's layout
    r = subprocess.run([exe, f"-w{w}", f"-p{pp}", *extra], cwd=run, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "Done. Image written to test.bmp." in r.stdout, r.stderr[-500:]
    got = host.read_bmp(os.path.join(run, "test.bmp"))
    core.upload_scene(worlds(w).scene(W, H, lens=bool(extra)))
    core.render(W, H, pp, seed=5)
    want, _, _ = core.readback(True, False)
    assert got.shape == want.shape and np.array_equal(got, want)
