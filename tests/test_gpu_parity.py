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


@pytest.mark.parametrize("w,W,H,pp", [(2, 1280, 720, 4), (3, 1280, 720, 4)])
def test_native_resolution_image_parity_through_image_compare(core, worlds, tmp_path, w, W, H, pp):
    """BASELINE configs C2 and C3 at their NATIVE 1280x720 (reduced spp so that the reference's own render, on every host
    core, takes seconds): the reference renders the frame twice (mt19937 seeds A, B), the GPU renders it at the same
    spp, and the repo's own bin/image_compare -- the tool north_star names -- must report
      * RMSE(reference A, GPU) <= 1.1 x RMSE(reference A, reference B): the GPU image is as close to the reference as the
        reference is to itself;
      * per-pixel z = (a - b) / sqrt(2 V), V = the variance of a pixel's mean at this spp MEASURED over 32 further GPU renders:
        |z| > 3 no more often for (reference, GPU) than for (reference, reference) (x1.5 + 0.2 %) and |median z| < 0.05
        (the estimator is heavy-tailed: mean and rms of z belong to a few firefly pixels, for the reference against itself too);
      * no bias: |global z| of the whole-image mean < 3 for the reference against the GPU render and against the MEAN of the
        32 GPU renders, each side with its own variance (the reference's per-pixel sums of squares, the GPU's 32 renders).
    The same comparison at every BASELINE configuration, larger spp: tools/parity_report.py."""
    import sys
    from conftest import ROOT
    sys.path.insert(0, ROOT + "/tools")
    import parity_report as P
    from oracle import ref_parallel as RP
    from pathtracer_b200 import host
    A = RP.render(w, W, H, pp, 0, 7, 1)
    B = RP.render(w, W, H, pp, 0, 7, 2)
    spp, K = pp * pp, 32
    core.upload_scene(worlds(w).scene(W, H))
    core.render(W, H, pp, seed=21)
    gpu_ldr, a, _ = core.readback(True, True)
    gpu_ldr, g1 = gpu_ldr.copy(), a[..., :3].astype(np.float64) / spp
    s1 = np.zeros((H, W, 3), np.float64)
    s2 = np.zeros((H, W, 3), np.float64)
    for k in range(K):
        core.render(W, H, pp, seed=1001 + k)
        _, a, _ = core.readback(False, True)
        m = a[..., :3].astype(np.float64) / spp
        s1 += m
        s2 += m * m
    gm = s1 / K
    V = np.maximum(0.0, s2 / K - gm * gm) * K / (K - 1)
    f = lambda n: str(tmp_path / n)
    host.write_bmp(f("a.bmp"), A["ldr"]); host.write_bmp(f("b.bmp"), B["ldr"]); host.write_bmp(f("g.bmp"), gpu_ldr)
    P.write_hdr(f("a.hdr"), A["sum"], spp, V); P.write_hdr(f("b.hdr"), B["sum"], spp, V)
    P.write_hdr(f("g.hdr"), g1 * spp, spp, V); P.write_hdr(f("gk.hdr"), gm * (K * spp), K * spp, V / K)
    ldr_g, ldr_n = P.parse_ldr(P.image_compare(f("a.bmp"), f("g.bmp"))), P.parse_ldr(P.image_compare(f("a.bmp"), f("b.bmp")))
    hdr_g, hdr_n = P.parse_hdr(P.image_compare("--hdr", f("a.hdr"), f("g.hdr"))), P.parse_hdr(P.image_compare("--hdr", f("a.hdr"), f("b.hdr")))
    P.write_hdr(f("a_own.hdr"), A["sum"], spp, RP.variance_of_mean(A))
    gz_1 = P.parse_hdr(P.image_compare("--hdr", f("a_own.hdr"), f("g.hdr")))["global_z"]
    gz_k = P.parse_hdr(P.image_compare("--hdr", f("a_own.hdr"), f("gk.hdr")))["global_z"]
    print(dict(world=w, ldr_ref_gpu=ldr_g, ldr_ref_ref=ldr_n, hdr_ref_gpu=hdr_g, hdr_ref_ref=hdr_n, global_z=(gz_1, gz_k), ref_kind=A["kind"]))
    assert ldr_g["rmse"] <= 1.1 * ldr_n["rmse"], (ldr_g, ldr_n)
    assert ldr_g["similarity"] >= ldr_n["similarity"] - 0.3, (ldr_g, ldr_n)
    assert hdr_g["frac_z3"] <= 1.5 * hdr_n["frac_z3"] + 0.2, (hdr_g, hdr_n)
    assert abs(hdr_g["median_z"]) < 0.05, hdr_g
    assert abs(gz_1) < 3.0 and abs(gz_k) < 3.0, (gz_1, gz_k)


# ---- isolated taps of the shading helpers -------------------------------------------------
def _one_sphere_scene(material, just_cosine=False, light_emit=15.0):
    """A unit sphere of `material` at the origin under a uniform sky, plus the (required) light sphere: far away and
    black for the no-light cases, overhead and bright otherwise.  Every path is camera -> sphere -> one BRDF sample
    -> sky / light, so the radiance of a ray IS the BRDF weight of that branch times a constant.  (light_emit = 0: the light
    sphere is an ordinary grey surface that the cone sampler still aims at -- the mixture pdf with a light that adds nothing.)"""
    import ctypes as C
    from pathtracer_b200 import host
    mats = (T.Material * 3)()
    for m in mats:
        m.alpha, m.ior, m.roughness = 1.0, 1.5, 1.0
    mats[0].emit = T.V3(0.6, 0.7, 0.9)                         # the sky (material 0)
    mats[1].emit = T.V3(light_emit, light_emit, light_emit)
    for k, v in material.items():
        setattr(mats[2], k, T.V3(*v) if isinstance(v, tuple) else v)
    sph = (T.Sphere * 2)(T.Sphere(T.V3(0.0, 0.0, 6.0), 1.0, 1), T.Sphere(T.V3(0, 0, 0), 1.0, 2))
    s = T.Scene()
    s.struct_size = C.sizeof(T.Scene)
    s.flags = T.SCENE_ALL_MAPS | (T.SCENE_JUST_COSINE if just_cosine else 0)
    s.n_materials, s.materials = 3, C.cast(mats, C.c_void_p)
    s.n_spheres, s.spheres = 2, C.cast(sph, C.c_void_p)
    s.camera.use_pinhole = 1
    s.camera.pos, s.camera.target, s.camera.fov = T.V3(0, -4, 0.5), T.V3(0, 0, 0), 40.0
    s.camera.focal_distance, s.camera.aperture_radius = 5.0, 0.035
    host.define_camera(s.camera, 64, 64)
    s._keep = (mats, sph)
    return s


_TAPS = [
    ("smooth metal: mirror branch + SchlickMetal", dict(roughness=0.0, metalness=1.0, metal_color=(0.9, 0.6, 0.3)), False, 15.0),
    ("smooth dielectric: Schlick F0 from ior, diffuse kd = 1 - ks", dict(roughness=0.0, metalness=0.0, albedo=(0.8, 0.2, 0.1)), False, 15.0),
    ("rough metal 0.3: GGX half vector, Hammon G, Burley alpha", dict(roughness=0.3, metalness=1.0, metal_color=(0.95, 0.93, 0.88)), False, 15.0),
    ("rough metal 0.7", dict(roughness=0.7, metalness=1.0, metal_color=(0.5, 0.8, 0.4)), False, 15.0),
    ("half metal 0.5 / rough 0.5: both lobes", dict(roughness=0.5, metalness=0.5, metal_color=(0.9, 0.9, 0.2), albedo=(0.2, 0.3, 0.9)), False, 15.0),
    ("pure diffuse with the light: cosine + cone mixture pdf", dict(roughness=1.0, metalness=0.0, albedo=(0.7, 0.7, 0.7)), False, 15.0),
    ("pure diffuse, JUST_COSINE: cosine pdf only", dict(roughness=1.0, metalness=0.0, albedo=(0.7, 0.5, 0.3)), True, 15.0),
    ("pure diffuse, black light: the cone samples land on a non-emitter", dict(roughness=1.0, metalness=0.0, albedo=(0.3, 0.6, 0.9)), False, 0.0),
]


@pytest.mark.parametrize("name,material,just_cosine,light", _TAPS, ids=[t[0].split(":")[0] for t in _TAPS])
def test_brdf_branches_in_isolation_vs_oracle(core, name, material, just_cosine, light):
    """pt_trace_paths on crafted one-sphere scenes, ray by ray against the oracle on the SAME Philox streams: a path is
    camera -> sphere -> one sample of one BRDF branch -> sky or light, so a compensating pair of errors inside
    pt_shade_vertex (SchlickMetal, Hammon G / Burley alpha, GGX half vector, cosine hemisphere, cone sampling of the
    light, the mixture pdf, the orthonormal basis) cannot hide behind a whole-image average.  Tolerance 1e-5 relative
    wherever no branch flips (a coin at 0.5 +- 1 ulp, a grazing N.L sign); such flips must stay below 0.2 % of rays
    (0.8 % where the path has up to three shaded vertices), and must not move the mean."""
    sc = _one_sphere_scene(material, just_cosine, light)
    core.upload_scene(sc)
    rng = np.random.default_rng(5)
    n = 20000
    # rays from a shell around the sphere aimed at points inside it: every incidence angle, every side
    o = rng.normal(size=(n, 3)); o = (o / np.linalg.norm(o, axis=1, keepdims=True) * rng.uniform(1.5, 4.0, (n, 1))).astype(np.float32)
    tgt = rng.normal(size=(n, 3)); tgt = tgt / np.linalg.norm(tgt, axis=1, keepdims=True) * rng.uniform(0, 0.95, (n, 1))
    d = (tgt - o); d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    got = core.trace_paths(o, d, seed=77).astype(np.float64)
    want = O.raycast(sc, o, d, O.RNG_PHILOX, 77).astype(np.float64)
    hits = core.trace_rays(o, d)
    assert (hits["kind"] == 1).all() and (hits["index"] == 1).mean() > 0.85          # the test sphere (or the light in front of it)
    scale = np.maximum(np.abs(want).max(1), 1e-3)
    rel = np.abs(got - want).max(1) / scale
    close = rel < 1e-5 * 20          # 1e-5 relative per operation chain; ~20 dependent float ops between draw and weight
    # with the black light the cone samples land on an ordinary surface and the path goes on for up to two more vertices:
    # three times the chances of a flip per ray
    flips_allowed = 0.002 if light else 0.008
    assert close.mean() > 1 - flips_allowed, f"{name}: {100 * (1 - close.mean()):.3f} % of rays differ (max rel {rel.max():.3e})"
    assert np.median(rel) < 2e-6, np.median(rel)
    # the branch really ran: radiance is neither all-sky nor all-black, and the three channels separate where the material says so
    assert np.isfinite(got).all() and got.std() > 1e-3
    assert abs(got.mean() - want.mean()) < 2e-3 * max(1.0, abs(want.mean()))
