"""GPU parity tests proper: the CUDA core, called through the C-ABI, against (a) fixtures made by the
reference itself, (b) the C oracle on identical inputs / identical Philox streams.

Bars (BASELINE.json north_star):
  * primary-ray hit primitive ids agree on >= 99.99 % of pixels (grazing / tie cases excepted);
  * resolve (float -> BGRA8) is byte-exact;
  * images within Monte-Carlo error of the reference: per-pixel z = (mu_gpu - mu_ref)/se within 3 sigma
    for all but a small fraction, no global bias.
"""
import numpy as np
import pytest

from conftest import golden
from oracle import pt_oracle as O
from pathtracer_b200 import _ctypes_api as T

pytestmark = pytest.mark.gpu

WORLDS = [1, 2, 3, 4, 5]


def _ids_equal(a_kind, a_index, b_kind, b_index):
    return (np.asarray(a_kind, np.int32) == np.asarray(b_kind, np.int32)) & (np.asarray(a_index, np.int32) == np.asarray(b_index, np.int32))


@pytest.mark.parametrize("w", WORLDS)
def test_primary_hit_ids_vs_reference_fixture(core, worlds, w):
    """160x90 jitter-free primary rays: ids from the reference's own RayCastIntersect (tests/golden)."""
    g = golden(f"hits_w{w}.npz")
    core.upload_scene(worlds(w).scene(160, 90))
    h = core.primary_hits(160, 90)
    same = _ids_equal(h["kind"], h["index"], g["kind"], g["index"])
    assert same.mean() >= 0.9999, f"only {same.mean():.6f} of primary ids agree"
    hit = (g["kind"] != 0) & same
    rel = np.abs(h["t"][hit] - g["t"][hit]) / g["t"][hit]
    assert rel.max() < 2e-3 and np.median(rel) < 1e-5
    assert (h["mat"][same] == g["mat"][same]).all()
    assert np.abs(h["normal"][same] - g["normal"][same]).max() < 5e-2
    miss = (g["kind"] == 0) & same
    assert (h["t"][miss] == np.float32(3.402823466e38)).all() and (h["mat"][miss] == 0).all()


@pytest.mark.parametrize("w,W,H", [(1, 1280, 720), (2, 1280, 720), (3, 1280, 720), (4, 1920, 1080), (5, 1920, 1080)])
def test_primary_hit_ids_full_size_vs_oracle(core, worlds, w, W, H):
    """Same check at the BASELINE config resolutions, against the oracle (bit-exact with the reference)."""
    sc = worlds(w).scene(W, H)
    core.upload_scene(sc)
    h = core.primary_hits(W, H)
    o, d = O.primary_rays(sc, W, H)
    r = O.intersect(sc, o, d)
    same = _ids_equal(h["kind"], h["index"], r["kind"], r["index"])
    assert same.mean() >= 0.9999, f"{(~same).sum()} of {same.size} primary ids differ"


@pytest.mark.parametrize("w", WORLDS)
def test_trace_rays_random_rays_vs_oracle(core, worlds, w):
    """Incoherent rays from points around the scene (secondary-ray-like), ids vs the oracle."""
    sc = worlds(w).scene(64, 36)
    core.upload_scene(sc)
    rng = np.random.default_rng(100 + w)
    n = 60000
    cam = np.array([sc.camera.pos.x, sc.camera.pos.y, sc.camera.pos.z], np.float32)
    tgt = np.array([sc.camera.target.x, sc.camera.target.y, sc.camera.target.z], np.float32)
    scale = np.float32(np.linalg.norm(cam - tgt))
    o = (tgt + (rng.random((n, 3)) - 0.5) * scale * 0.8).astype(np.float32)
    o[:, 2] = np.abs(o[:, 2]) + np.float32(0.05)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    h = core.trace_rays(o, d)
    r = O.intersect(sc, o, d)
    same = _ids_equal(h["kind"], h["index"], r["kind"], r["index"])
    assert same.mean() >= 0.9995, f"{(~same).sum()} of {n} ids differ"


def test_resolve_byte_exact(core):
    g = golden("common.npz")
    assert np.array_equal(core.resolve_colors(g["resolve_in"]), g["resolve_out"])
    rng = np.random.default_rng(5)
    cols = np.concatenate([rng.random((400000, 3)) ** 4 * 8, rng.random((100000, 3)) * 0.004,
                           np.array([[np.nan, np.inf, -np.inf], [-1, 0, 1e30]])]).astype(np.float32)
    assert np.array_equal(core.resolve_colors(cols), O.resolve(cols))


@pytest.mark.parametrize("w", WORLDS)
def test_paths_same_streams_vs_oracle(core, worlds, w):
    """Radiance of the same rays with the same Philox streams: equal up to float rounding except where
    a rounding difference flips a branch (hit/miss at a silhouette, a coin at 0.5)."""
    sc = worlds(w).scene(160, 90)
    core.upload_scene(sc)
    o, d = O.primary_rays(sc, 160, 90)
    a = O.raycast(sc, o, d, O.RNG_PHILOX, 21)
    b = core.trace_paths(o, d, 21)
    fin = np.isfinite(a).all(1) & np.isfinite(b).all(1)
    assert fin.mean() > 0.999
    rel = np.abs(a - b).max(1) / np.maximum(1e-3, np.abs(a).max(1))
    close = (rel[fin] < 1e-3).mean()
    assert close > (0.97 if w in (1, 4) else 0.995), close
    assert np.allclose(a[fin].mean(0), b[fin].mean(0), rtol=5e-3, atol=1e-4)


@pytest.mark.parametrize("w,lens,flags", [(1, False, 7), (2, False, 7), (3, False, 7), (4, False, 7), (5, False, 7),
                                          (3, True, 7), (5, True, 7), (1, False, 0), (1, False, T.SCENE_ROUGHNESS_MAPS)])
def test_render_same_streams_vs_oracle(core, worlds, w, lens, flags):
    """pt_render vs the oracle's render on identical streams: pixel means, sample counts, LDR bytes."""
    W, H, pp = 80, 45, 4
    sc = worlds(w).scene(W, H, lens=lens, tex_flags=flags)
    core.upload_scene(sc)
    core.render(W, H, pp, seed=9)
    bgra, acc, st = core.readback(True, True)
    orc = O.render(sc, W, H, pp, O.RNG_PHILOX, 9)
    assert (acc[..., 3] == pp * pp).all() and st.samples == W * H * pp * pp
    gm = acc[..., :3] / (pp * pp)
    om = orc["sum"] / (pp * pp)
    rel = np.abs(gm - om).max(2) / np.maximum(1e-2, np.abs(om).max(2))
    assert (rel < 2e-3).mean() > (0.93 if w in (1, 4) else 0.99), (rel < 2e-3).mean()
    assert np.allclose(gm.mean((0, 1)), om.mean((0, 1)), rtol=1e-2)
    assert (bgra == orc["ldr"]).mean() > (0.92 if w in (1, 4) else 0.985)
    assert (bgra >> 24 == 255).all()


def _zstats(mu_gpu, se_gpu2, s, s2, n):
    mu = s / n
    var = np.maximum(s2 / n - mu ** 2, 0) * n / (n - 1)
    se = np.sqrt(var / n + se_gpu2)
    ok = se > 1e-9
    z = (mu_gpu - mu)[ok] / se[ok]
    npx = mu.shape[0] * mu.shape[1]
    g_se = np.sqrt((var / n + se_gpu2).sum((0, 1))) / npx
    g_z = (mu_gpu.mean((0, 1)) - mu.mean((0, 1))) / np.maximum(g_se, 1e-12)
    return dict(frac3=float((np.abs(z) > 3).mean()), med=float(np.median(z)), g_z=g_z, ok=ok, mu=mu)


@pytest.mark.parametrize("name,w,lens,flags", [
    ("hdr_w1.npz", 1, False, 7), ("hdr_w2.npz", 2, False, 7), ("hdr_w3.npz", 3, False, 7),
    ("hdr_w4.npz", 4, False, 7), ("hdr_w5.npz", 5, False, 7),
    ("hdr_w3_lens.npz", 3, True, 7), ("hdr_w1_nmr.npz", 1, False, 0)])
def test_image_within_monte_carlo_error_of_reference(core, worlds, name, w, lens, flags):
    """Statistical parity against the REFERENCE's own render (fixture: per-pixel sum and sum of squares of
    its mt19937 samples, tools/make_golden.py).  The GPU renders the same pixels with 16x the samples from
    8 seeds (so its own error is small and measured); z = (mu_gpu - mu_ref)/sqrt(se_ref^2 + se_gpu^2).

    The estimator is heavy-tailed (sun radiance 15 x weights 2/px; 64..576 spp), so z is NOT N(0,1) even
    for the reference against itself.  The test therefore calibrates against the null: the same statistic
    with two fresh draws of the reference's estimator (the oracle, bit-exact with the reference, mt19937
    seeds 99/100, same spp) in place of the fixture.  The fixture must be indistinguishable from them:
      * fraction of pixel-channels beyond 3 sigma no larger than the null's (x1.6 + 1 %);
      * median z within 0.2 of the null's; global image mean within 5 standard errors per channel."""
    g = golden(name)
    pp = int(g["pp"])
    n = pp * pp
    W, H = 48, 27
    sc = worlds(w).scene(W, H, lens=lens, tex_flags=flags)
    core.upload_scene(sc)
    means = []
    for seed in range(8):
        core.render(W, H, pp * 4, seed=1000 + seed)
        _, acc, _ = core.readback(False, True)
        means.append(acc[..., :3].astype(np.float64) / acc[..., 3:4])
    means = np.stack(means)
    mu_gpu = means.mean(0)
    se_gpu2 = means.var(0, ddof=1) / len(means)
    ref = _zstats(mu_gpu, se_gpu2, g["sum"], g["sumsq"], n)
    nulls = []
    for seed in (99, 100):
        O.seed_mt(seed)
        r = O.render(sc, W, H, pp, O.RNG_MT19937)
        nulls.append(_zstats(mu_gpu, se_gpu2, r["sum"], r["sumsq"], n))
    null_frac3 = max(x["frac3"] for x in nulls)
    null_med = np.mean([x["med"] for x in nulls])
    flat = ~ref["ok"]                     # pixels the reference saw with zero variance: means must be equal
    if flat.any():
        assert np.allclose(mu_gpu[flat], ref["mu"][flat], rtol=2e-3, atol=2e-3)
    assert ref["frac3"] <= 1.6 * null_frac3 + 0.01, (ref["frac3"], null_frac3)
    assert abs(ref["med"] - null_med) < 0.2, (ref["med"], null_med)
    assert (np.abs(ref["g_z"]) < 5).all(), ref["g_z"]
    assert np.allclose(mu_gpu.mean((0, 1)), ref["mu"].mean((0, 1)), rtol=0.08)


@pytest.mark.parametrize("seed,just_cosine,lens", [(11, False, False), (12, True, False), (13, False, True)])
def test_synthetic_scene_with_boxes_vs_oracle(core, seed, just_cosine, lens):
    """The primitives no built-in world holds -- axis-aligned boxes (main.cpp:529-553) -- next to quads, spheres and a
    plane in one scene: closest-hit ids on rays from outside and inside the boxes, then whole renders on identical
    Philox streams, against the oracle (itself bit-exact with the reference's box function, tests/golden/aabb.npz)."""
    from synth import SynthScene
    sc = SynthScene(seed=seed, just_cosine=just_cosine)
    if lens:
        sc.scene.camera.use_pinhole = 0
    core.upload_scene(sc.scene)
    o, d = sc.random_rays(40000, seed)
    h = core.trace_rays(o, d)
    r = O.intersect(sc.scene, o, d)
    same = _ids_equal(h["kind"], h["index"], r["kind"], r["index"])
    assert same.mean() >= 0.9995, f"{(~same).sum()} of {len(o)} ids differ"
    assert (r["kind"] == 5).mean() > 0.2                       # boxes are what most rays hit
    hit = same & (r["kind"] != 0)
    assert np.abs(h["t"][hit] - r["t"][hit]).max() <= 2e-3 * r["t"][hit].max()
    assert (h["mat"][same] == r["mat"][same]).all()
    assert np.abs(h["normal"][same] - r["normal"][same]).max() < 1e-3
    W, H, pp = sc.width, sc.height, 4
    core.render(W, H, pp, seed=9)
    bgra, acc, st = core.readback(True, True)
    orc = O.render(sc.scene, W, H, pp, O.RNG_PHILOX, 9)
    assert (acc[..., 3] == pp * pp).all()
    gm, om = acc[..., :3] / (pp * pp), orc["sum"] / (pp * pp)
    rel = np.abs(gm - om).max(2) / np.maximum(1e-2, np.abs(om).max(2))
    assert (rel < 2e-3).mean() > 0.97, (rel < 2e-3).mean()
    assert np.allclose(gm.mean((0, 1)), om.mean((0, 1)), rtol=1e-2)
    assert (bgra == orc["ldr"]).mean() > 0.96


def _legacy_similarity(a, b):
    """The reference's ImageCompare metric on BGRA words (src/image_compare.c:79-134; red ignored, alpha twice)."""
    ab, bb = a.view(np.uint8).reshape(a.shape + (4,)).astype(np.float64), b.view(np.uint8).reshape(b.shape + (4,)).astype(np.float64)
    d = (ab - bb) / 255.0
    return float((1.0 - np.sqrt(2 * d[..., 3] ** 2 + d[..., 0] ** 2 + d[..., 1] ** 2)).mean() * 100.0)


def _rmse(a, b):
    ab, bb = a.view(np.uint8).reshape(a.shape + (4,))[..., :3].astype(np.float64), b.view(np.uint8).reshape(b.shape + (4,))[..., :3].astype(np.float64)
    return float(np.sqrt(((ab - bb) ** 2).mean()))


@pytest.mark.parametrize("w,pp", [(3, 8), (5, 8)])
def test_ldr_image_within_the_references_own_noise_floor(core, worlds, w, pp):
    """the 8-bit image the user sees.  Two independent renders by the reference's estimator (the oracle
    in mt19937 mode, bit-exact with the reference) differ by Monte-Carlo noise; the GPU image must be as close to
    either of them as they are to each other, in the reference's own "Percentage Similarity" and in RMSE."""
    W, H = 160, 90
    sc = worlds(w).scene(W, H)
    refs = []
    for seed in (31, 32):
        O.seed_mt(seed)
        refs.append(O.render(sc, W, H, pp, O.RNG_MT19937)["ldr"])
    core.upload_scene(sc)
    core.render(W, H, pp, seed=5)
    gpu, _, _ = core.readback(True, False)
    floor_sim, floor_rmse = _legacy_similarity(refs[0], refs[1]), _rmse(refs[0], refs[1])
    for r in refs:
        assert _legacy_similarity(gpu, r) >= floor_sim - 0.4, (_legacy_similarity(gpu, r), floor_sim)
        assert _rmse(gpu, r) <= floor_rmse * 1.08 + 0.2, (_rmse(gpu, r), floor_rmse)
