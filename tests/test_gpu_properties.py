"""Size-independent properties of the CUDA path (checked up to BASELINE.json's full sizes), edge cases
and the error behaviour of the C-ABI."""
import ctypes as C

import numpy as np
import pytest

from oracle import pt_oracle as O
from pathtracer_b200 import _ctypes_api as T
from pathtracer_b200 import render

pytestmark = pytest.mark.gpu


def _frame(core, W, H, pp, **kw):
    core.render(W, H, pp, **kw)
    b, a, st = core.readback(True, True)
    return b.copy(), a.copy(), st


def test_deterministic_and_seed_dependent(core, worlds):
    core.upload_scene(worlds(5).scene(200, 120))
    b1, a1, _ = _frame(core, 200, 120, 4, seed=3)
    b2, a2, _ = _frame(core, 200, 120, 4, seed=3)
    b3, a3, _ = _frame(core, 200, 120, 4, seed=4)
    assert np.array_equal(a1, a2) and np.array_equal(b1, b2)
    assert not np.array_equal(a1, a3)


@pytest.mark.parametrize("G", [2, 3, 8])
def test_tile_sharding_is_bit_identical(core, worlds, G):
    """Counter RNG keyed by (pixel, sample): the union of the ranks' tiles equals the 1-GPU frame, bit for bit."""
    W, H, pp = 300, 170, 4          # not multiples of 32: clipped edge tiles
    core.upload_scene(worlds(5).scene(W, H))
    full_b, full_a, _ = _frame(core, W, H, pp, seed=5)
    acc = np.zeros_like(full_a)
    covered = np.zeros((H, W), np.int32)
    for r in range(G):
        b, a, st = _frame(core, W, H, pp, seed=5, tile_rank=r, tile_count=G)
        mine = a[..., 3] > 0
        covered += mine
        acc[mine] = a[mine]
        for t in render.tile_ids(W, H, r, G):
            tx, ty = (t % ((W + 31) // 32)) * 32, (t // ((W + 31) // 32)) * 32
            assert mine[ty:ty + 32, tx:tx + 32].all()
        assert st.samples == mine.sum() * pp * pp
    assert (covered == 1).all()
    assert np.array_equal(acc, full_a)


def test_tile_pack_unpack_roundtrip(core, worlds):
    import torch
    W, H, pp, G = 300, 170, 2, 3
    core.upload_scene(worlds(3).scene(W, H))
    full_b, _, _ = _frame(core, W, H, pp, seed=5)
    packs = []
    for r in range(G):
        _frame(core, W, H, pp, seed=5, tile_rank=r, tile_count=G)
        from pathtracer_b200 import dist as D
        n = D.slots_per_rank(W, H, G)
        buf = torch.zeros(n * 1024, dtype=torch.int32, device="cuda")
        assert core.pack_tiles(buf.data_ptr(), r, G) == n
        core.synchronize()
        packs.append(buf)
    _frame(core, W, H, 0, seed=5)          # clears accum; bgra of a zero-sample frame
    for r in range(G):
        core.unpack_tiles(packs[r].data_ptr(), r, G)
    core.synchronize()
    _, bdev = core.device_buffers()
    got = torch.as_tensor(bdev, device="cuda").cpu().numpy().view(np.uint32).reshape(H, W)
    assert np.array_equal(got, full_b)


def test_sample_sharding_sums_to_full_frame(core, worlds):
    W, H, pp = 160, 90, 6
    core.upload_scene(worlds(3).scene(W, H))
    _, full, _ = _frame(core, W, H, pp, seed=8)
    parts = np.zeros_like(full, dtype=np.float64)
    bounds = [0, 9, 18, 27, 36]
    for lo, hi in zip(bounds[:-1], bounds[1:]):
        _, a, st = _frame(core, W, H, pp, seed=8, sample_begin=lo, sample_end=hi)
        assert (a[..., 3] == hi - lo).all() and st.samples == W * H * (hi - lo)
        parts += a
    assert (parts[..., 3] == pp * pp).all()
    assert np.allclose(parts[..., :3], full[..., :3], rtol=2e-5, atol=1e-5)
    # accumulate-without-clear gives the same as the sum
    core.render(W, H, pp, seed=8, sample_begin=0, sample_end=18, clear=True)
    core.render(W, H, pp, seed=8, sample_begin=18, sample_end=36, clear=False)
    _, a2, st = core.readback(False, True)
    assert st.samples == W * H * 36 and np.allclose(a2[..., :3], full[..., :3], rtol=2e-5, atol=1e-5)


@pytest.mark.parametrize("lanes", [0, 1, 2, 3, 4, 5])
def test_launch_shape_does_not_change_the_samples(worlds, lanes):
    """S sample-lanes x P pixels per warp is a scheduling choice only: the kernel adds a pixel's samples in a
    canonical order (32 logical lanes, xor-butterfly), so every launch shape gives BIT-identical sums -- which is
    what lets small multi-GPU shares use more lanes per pixel than a full frame and still reproduce it exactly."""
    W, H, pp = 100, 70, 6
    rc = render.RenderCore(0)
    rc.set_tuning(sample_lanes_log2=lanes)
    rc.upload_scene(worlds(5).scene(W, H))
    rc.render(W, H, pp, seed=2)
    _, a, _ = rc.readback(False, True)
    rc2 = render.RenderCore(0)
    rc2.upload_scene(worlds(5).scene(W, H))
    rc2.render(W, H, pp, seed=2)
    _, b, _ = rc2.readback(False, True)
    assert np.array_equal(a, b)
    rc.close(); rc2.close()

@pytest.mark.parametrize("w,lens", [(1, False), (2, False), (3, False), (4, False), (5, False), (3, True), (5, True)])
def test_wavefront_kernel_renders_the_same_samples(worlds, w, lens):
    """PT_KERNEL_WAVEFRONT (paths pooled in shared memory, stage lists per warp) and the megakernel are two
    schedules of the same samples: identical sample counts and ray counts, pixel sums equal up to float
    summation order, LDR bytes equal except where a sum sits on a rounding boundary."""
    W, H, pp = 150, 90, 6
    sc = worlds(w).scene(W, H, lens=lens)
    out = []
    for kernel in (T.KERNEL_MEGA, T.KERNEL_WAVEFRONT):
        rc = render.RenderCore(0, kernel=kernel)
        rc.upload_scene(sc)
        rc.render(W, H, pp, seed=3)
        out.append(rc.readback(True, True))
        rc.close()
    (b0, a0, s0), (b1, a1, s1) = out
    assert np.array_equal(a0[..., 3], a1[..., 3]) and s0.rays == s1.rays and s0.nan_retries == s1.nan_retries
    assert np.allclose(a0[..., :3], a1[..., :3], rtol=2e-5, atol=1e-5)
    assert (b0 == b1).mean() > 0.999


@pytest.mark.parametrize("w,lens", [(5, False), (5, True), (2, False), (4, False)])
def test_cooperative_sparse_walks_do_not_change_the_samples(worlds, w, lens):
    """The PT_F_COOP kernel family traces the walks that few lanes of a warp enter one ray at a time with all 32
    lanes over the cluster cut of the tree (pt_cluster.h, pt_coop_trace): the same candidates, the same
    (t, id) minimum -- bit-identical sums and ray counts for any threshold, on triangle and sphere trees.
    (CPU side of the argument: tests/test_clusters.py.)"""
    W, H, pp = 200, 120, 5
    sc = worlds(w).scene(W, H, lens=lens)
    res = []
    for coop in (0, -1, 1, 5, 32):            # off, the default (12 for triangles, off for spheres), thresholds
        rc = render.RenderCore(0)
        rc.set_tuning(coop_max=coop)                     # before the upload: the cluster cut is built there
        rc.upload_scene(sc)
        rc.render(W, H, pp, seed=11)
        _, a, st = rc.readback(False, True)
        res.append((a.copy(), st.rays, st.nan_retries))
        rc.close()
    for a, rays, nan in res[1:]:
        assert np.array_equal(a.view(np.uint32), res[0][0].view(np.uint32))
        assert rays == res[0][1] and nan == res[0][2]


@pytest.mark.parametrize("batch", [1, 8, 32])
def test_shade_deferral_does_not_change_the_samples(worlds, batch):
    """Parking lanes until `shade_batch` of them need a BRDF sample is scheduling only: bit-identical sums."""
    W, H, pp = 120, 80, 6
    res = []
    for b in (0, batch):
        for w in (1, 5):
            rc = render.RenderCore(0)
            rc.set_tuning(shade_batch=b)
            rc.upload_scene(worlds(w).scene(W, H))
            rc.render(W, H, pp, seed=4)
            res.append(rc.readback(False, True)[1])
            rc.close()
    assert np.array_equal(res[0], res[2]) and np.array_equal(res[1], res[3])

@pytest.mark.parametrize("w,W,H", [(2, 320, 180), (4, 320, 180), (5, 480, 270)])
def test_device_lbvh_gives_the_same_hits_and_image(worlds, w, W, H):
    """PT_BVH_DEVICE_LBVH (Morton -> radix sort -> Karras -> refit on the GPU) is another tree over the same
    primitives with the same accept rule: identical primary ids and t, and -- since every sample then follows
    the same path -- a bit-identical accumulator."""
    sc = worlds(w).scene(W, H)
    out = []
    for bvh in (T.BVH_HOST_SAH, T.BVH_DEVICE_LBVH):
        rc = render.RenderCore(0, bvh=bvh)
        rc.upload_scene(sc)
        hits = rc.primary_hits(W, H)
        rc.render(W, H, 4, seed=6)
        out.append((hits, rc.readback(True, True)))
        rc.close()
    (h0, (b0, a0, s0)), (h1, (b1, a1, s1)) = out
    assert np.array_equal(h0["kind"], h1["kind"]) and np.array_equal(h0["index"], h1["index"]) and np.array_equal(h0["t"], h1["t"])
    assert s0.rays == s1.rays
    assert np.array_equal(a0, a1) and np.array_equal(b0, b1)
    assert s1.bvh_nodes > 0 and s1.bvh_prims == s0.bvh_prims

@pytest.mark.parametrize("G", [2, 3])
def test_single_process_peer_gather_and_reduce(worlds, G):
    """pt_gather_tiles_from / pt_add_accum_from (what `pathtracer --gpus N` uses): G contexts render their shards,
    context 0 pulls the rest by peer copies.  Tiles: bit-identical to the unsharded frame.  Samples: equal up to
    float summation order.  (Contexts share device 0 here; the copies are the same calls across devices.)"""
    W, H, pp = 200, 130, 4
    sc = worlds(5).scene(W, H)
    one = render.RenderCore(0)
    one.upload_scene(sc)
    one.render(W, H, pp, seed=12)
    b_ref, a_ref, _ = one.readback(True, True)
    cores = [render.RenderCore(0) for _ in range(G)]
    for c in cores:
        c.upload_scene(sc)
    for r, c in enumerate(cores):
        c.render(W, H, pp, seed=12, tile_rank=r, tile_count=G)
    for r in range(1, G):
        cores[0].gather_tiles_from(cores[r], r, G)
    b, _, _ = cores[0].readback(True, False)
    assert np.array_equal(b, b_ref)
    spp = pp * pp
    for r, c in enumerate(cores):
        c.render(W, H, pp, seed=12, sample_begin=r * spp // G, sample_end=(r + 1) * spp // G)
    for r in range(1, G):
        cores[0].add_accum_from(cores[r])
    b2, a2, st = cores[0].readback(True, True)
    assert (a2[..., 3] == spp).all() and st.samples == W * H * spp
    assert np.allclose(a2[..., :3], a_ref[..., :3], rtol=2e-5, atol=1e-5)
    assert (b2 == b_ref).mean() > 0.999
    for c in cores + [one]:
        c.close()


@pytest.mark.parametrize("W,H,pp", [(1, 1, 3), (33, 31, 1), (31, 65, 2), (64, 64, 5)])
def test_ragged_sizes_vs_oracle(core, worlds, W, H, pp):
    sc = worlds(3).scene(W, H)
    core.upload_scene(sc)
    b, a, st = _frame(core, W, H, pp, seed=1)
    assert a.shape == (H, W, 4) and (a[..., 3] == pp * pp).all()
    orc = O.render(sc, W, H, pp, O.RNG_PHILOX, 1)
    om = orc["sum"] / (pp * pp)
    gm = a[..., :3] / (pp * pp)
    rel = np.abs(gm - om).max(2) / np.maximum(1e-2, np.abs(om).max(2))
    assert (rel < 2e-3).mean() >= 0.97


def test_zero_samples_frame(core, worlds):
    """-p0 is legal in the reference (clamp to [0,1000], main.cpp:2171): no samples, colour 0 -> black, A=255."""
    core.upload_scene(worlds(2).scene(40, 30))
    b, a, st = _frame(core, 40, 30, 0)
    assert (a == 0).all() and st.samples == 0 and (b == 0xFF000000).all()


def test_max_pp_is_accepted_and_more_is_rejected(core, worlds):
    core.upload_scene(worlds(3).scene(2, 2))
    b, a, st = _frame(core, 2, 2, 1000, sample_begin=0, sample_end=64)     # RAYS_PER_PIXEL_MAX, main.cpp:89
    assert (a[..., 3] == 64).all()
    with pytest.raises(render.PtError):
        core.render(2, 2, 1001)


def test_empty_primitive_lists(core):
    """A world with nothing but the sky and the (required) light sphere far away: every ray misses."""
    from pathtracer_b200 import host
    mats = (T.Material * 2)()
    for m in mats:
        m.alpha, m.ior, m.roughness = 1, 1, 1
    mats[0].emit = T.V3(0.25, 0.5, 0.75)
    mats[1].emit = T.V3(15, 15, 15)
    sph = (T.Sphere * 1)(T.Sphere(T.V3(0, -5000, 0), 1.0, 1))
    s = T.Scene()
    s.struct_size = C.sizeof(T.Scene)
    s.n_materials, s.materials = 2, C.cast(mats, C.c_void_p)
    s.n_spheres, s.spheres = 1, C.cast(sph, C.c_void_p)
    s.camera.use_pinhole = 1
    s.camera.pos, s.camera.target, s.camera.fov = T.V3(0, -10, 1), T.V3(0, 0, 0), 45.0
    s.camera.focal_distance, s.camera.aperture_radius = 5.0, 0.035
    host.define_camera(s.camera, 32, 16)
    core.upload_scene(s)
    h = core.primary_hits(32, 16)
    assert (h["kind"] == 0).all() and (h["index"] == -1).all()
    b, a, _ = _frame(core, 32, 16, 2)
    assert np.allclose(a[..., :3] / 4, [0.25, 0.5, 0.75])
    assert (b == O.resolve(np.array([[0.25, 0.5, 0.75]], np.float32))[0]).all()


def test_error_codes(core, worlds):
    L = render.lib()
    sc = worlds(3).scene(16, 16)
    bad = T.Scene.from_buffer_copy(sc)
    bad.struct_size = 4
    assert L.pt_upload_scene(core._h, C.byref(bad)) == -1 and b"struct_size" in L.pt_last_error(core._h)
    bad = T.Scene.from_buffer_copy(sc)
    bad.n_materials = 0
    assert L.pt_upload_scene(core._h, C.byref(bad)) == -1
    sph = np.frombuffer((C.c_uint8 * (sc.n_spheres * 20)).from_address(sc.spheres), T.SPHERE_DTYPE).copy()
    sph["mat"][0] = 999
    bad = T.Scene.from_buffer_copy(sc)
    bad.spheres = sph.ctypes.data
    assert L.pt_upload_scene(core._h, C.byref(bad)) == -1 and b"material" in L.pt_last_error(core._h)
    fresh = render.RenderCore(0)
    with pytest.raises(render.PtError) as e:
        fresh.render(16, 16, 2)
    assert e.value.code == -4
    p = T.RenderParams(C.sizeof(T.RenderParams), 16, 16, 2, 1, 0, 0, 0, 1, 1)
    assert L.pt_readback(fresh._h, None, None, None) == -5
    fresh.upload_scene(sc)
    p.tile_rank, p.tile_count = 3, 2
    assert L.pt_render(fresh._h, C.byref(p)) == -1
    p.tile_rank, p.tile_count, p.sample_begin, p.sample_end = 0, 1, 3, 2
    assert L.pt_render(fresh._h, C.byref(p)) == -1
    p.sample_begin, p.sample_end, p.width = 0, 0, -5
    assert L.pt_render(fresh._h, C.byref(p)) == -1
    cfg = T.Config(C.sizeof(T.Config), 4096, 0, 0, None)
    h = C.c_void_p()
    assert L.pt_create(C.byref(cfg), C.byref(h)) == -1
    fresh.close()


def test_headline_config_properties(core, worlds):
    """w5 1920x1080 -p32 (the metric's configuration): sample counts, energy vs a low-spp oracle strip,
    tile sharding at full size bit-identical on a strip, no NaN in the accumulation buffer."""
    W, H, pp = 1920, 1080, 32
    sc = worlds(5).scene(W, H)
    core.upload_scene(sc)
    b, a, st = _frame(core, W, H, pp, seed=1)
    assert st.samples == W * H * pp * pp and (a[..., 3] == pp * pp).all()
    assert np.isfinite(a).all()
    assert 1.35 < st.rays / st.samples < 1.50          # the reference measures 1.418 rays per sample
    mean = a[..., :3] / (pp * pp)
    # a 64x4 strip through Mario against the oracle at 64 spp (different sample count -> statistical)
    x0, y0 = 930, 500
    orc = O.render(sc, W, H, 8, O.RNG_PHILOX, 77, window=(x0, y0, 64, 4))
    om = orc["sum"] / 64
    assert np.allclose(mean[y0:y0 + 4, x0:x0 + 64].mean((0, 1)), om.mean((0, 1)), rtol=0.05, atol=2e-3)
    # two of eight tile ranks reproduce their pixels bit-identically
    for r in (0, 5):
        _, ar, _ = _frame(core, W, H, pp, seed=1, tile_rank=r, tile_count=8)
        mine = ar[..., 3] > 0
        assert np.array_equal(ar[mine], a[mine]) and abs(mine.mean() - 1 / 8) < 0.01


def test_resolve_pack_and_one_kernel_unpack_equal_the_full_frame(core, worlds):
    """The multi-GPU tile exchange as it runs now: every rank resolves ITS tiles straight into the tile-major pack
    (pt_resolve_pack_tiles), the packs sit back to back in one gathered buffer, and one kernel scatters all of them
    (pt_unpack_gathered).  Byte-identical to the unsharded frame; also to the host mirror used by the gloo tests."""
    import torch
    from pathtracer_b200 import dist as D
    W, H, pp, G = 300, 170, 3, 3
    core.upload_scene(worlds(5).scene(W, H))
    full_b, _, _ = _frame(core, W, H, pp, seed=9)
    n_max = D.max_tiles_per_rank(W, H, G) * 1024
    gathered = torch.zeros((G, n_max), dtype=torch.int32, device="cuda")
    for r in range(G):
        core.render(W, H, pp, seed=9, tile_rank=r, tile_count=G)
        n = core.resolve_pack_tiles(gathered[r].data_ptr(), r, G)
        assert n == D.slots_per_rank(W, H, G)
        core.synchronize()
    core.render(W, H, 0, seed=9)                       # empty accumulator: the frame must come from the packs alone
    core.unpack_gathered(gathered.data_ptr(), G, n_max)
    b, _, _ = core.readback(True, False)               # marked resolved: readback must not resolve the empty accumulator
    assert np.array_equal(b, full_b)
    host = D.unpack_gathered_host(gathered.cpu().numpy().view(np.uint32), W, H, G)
    assert np.array_equal(host, full_b)
    with pytest.raises(render.PtError):
        core.unpack_gathered(gathered.data_ptr(), G, 1024)      # stride shorter than a rank's pack


def test_resolve_range_matches_resolve(core, worlds):
    """pt_resolve_range (a rank's share of the pixels after the reduce-scatter) is the same arithmetic as pt_resolve."""
    import torch
    W, H, pp = 160, 90, 4
    core.upload_scene(worlds(3).scene(W, H))
    b, a, _ = _frame(core, W, H, pp, seed=2)
    adev, _ = core.device_buffers()
    acc = torch.as_tensor(adev, device="cuda").view(-1).clone()
    out = torch.zeros(W * H, dtype=torch.int32, device="cuda")
    lo, hi = 1000, 9000
    core.resolve_range(acc[lo * 4:].data_ptr(), out[lo:].data_ptr(), hi - lo)
    core.synchronize()
    got = out.cpu().numpy().view(np.uint32)
    assert np.array_equal(got[lo:hi], b.reshape(-1)[lo:hi]) and (got[:lo] == 0).all() and (got[hi:] == 0).all()


def test_keep_textures_flag_skips_the_texel_upload_only(worlds):
    """PT_SCENE_KEEP_TEXTURES: same (pointer, width, height) as the previous upload -> the device texels are kept
    (pt_stats.texture_bytes_uploaded == 0) and the image is bit-identical; without the flag everything is re-sent."""
    W, H, pp = 96, 54, 3
    sc = worlds(1).scene(W, H)
    rc = render.RenderCore(0)
    rc.upload_scene(sc)
    rc.render(W, H, pp, seed=3)
    b0, a0, st0 = rc.readback(True, True)
    tex_bytes = sum(sc.textures[i].width * sc.textures[i].height * 4 for i in range(4) if sc.textures[i].rgba8)
    assert st0.texture_bytes_uploaded == tex_bytes > 0
    keep = T.Scene.from_buffer_copy(sc)
    keep.flags |= T.SCENE_KEEP_TEXTURES
    rc.upload_scene(keep)
    rc.render(W, H, pp, seed=3)
    b1, a1, st1 = rc.readback(True, True)
    assert st1.texture_bytes_uploaded == 0 and np.array_equal(a0, a1) and np.array_equal(b0, b1)
    rc.upload_scene(sc)                                 # no flag: sent again
    rc.render(W, H, pp, seed=3)
    _, a2, st2 = rc.readback(False, True)
    assert st2.texture_bytes_uploaded == tex_bytes and np.array_equal(a0, a2)
    # a different scene in between, then the textured one again with the flag: the slots' sizes still match, the pointers do too
    rc.upload_scene(worlds(3).scene(W, H))
    rc.upload_scene(keep)
    rc.render(W, H, pp, seed=3)
    _, a3, st3 = rc.readback(False, True)
    assert np.array_equal(a0, a3)
    rc.close()


def test_scene_reupload_is_stable(worlds):
    """The scene arena (one pinned staging buffer, one device buffer, one copy) is reused and grown across uploads:
    small scene -> large scene -> small scene gives the same images as fresh contexts."""
    W, H, pp = 80, 48, 3
    fresh = {}
    for w in (3, 5, 4):
        rc = render.RenderCore(0)
        rc.upload_scene(worlds(w).scene(W, H))
        rc.render(W, H, pp, seed=6)
        fresh[w] = rc.readback(False, True)[1].copy()
        rc.close()
    rc = render.RenderCore(0)
    for w in (3, 5, 4, 3, 5):
        rc.upload_scene(worlds(w).scene(W, H))
        rc.render(W, H, pp, seed=6)
        assert np.array_equal(rc.readback(False, True)[1], fresh[w]), w
    rc.close()


def test_tuning_is_validated_and_the_library_reads_no_environment(worlds, monkeypatch):
    L = render.lib()
    rc = render.RenderCore(0)
    t = T.Tuning(C.sizeof(T.Tuning), -1, 0, -1, 0, 0, -1, 0)
    assert L.pt_set_tuning(rc._h, C.byref(t)) == 0
    for field, bad in [("sample_lanes_log2", 6), ("shade_batch", 33), ("coop_max", 40), ("tri_leaf", 65), ("wf_pixels_log2", 11)]:
        u = T.Tuning.from_buffer_copy(t)
        setattr(u, field, bad)
        assert L.pt_set_tuning(rc._h, C.byref(u)) == -1, field
    t.struct_size = 8
    assert L.pt_set_tuning(rc._h, C.byref(t)) == -1
    cfg = T.Config(4, 0, 0, 0, None)
    h = C.c_void_p()
    assert L.pt_create(C.byref(cfg), C.byref(h)) == -1 and b"struct_size" in L.pt_last_error(None)
    rc.close()
    # an environment variable that the round-1 library obeyed must change nothing: same kernel, same launch shape
    W, H, pp = 64, 40, 3
    out = []
    for env in (None, "wavefront"):
        if env:
            monkeypatch.setenv("PT_KERNEL", env)
            monkeypatch.setenv("PT_SAMPLE_LANES_LOG2", "0")
        rc = render.RenderCore(0)
        rc.upload_scene(worlds(5).scene(W, H))
        rc.render(W, H, pp, seed=1)
        out.append(rc.readback(False, True))
        rc.close()
    assert np.array_equal(out[0][1], out[1][1]) and out[0][2].kernel_launches == out[1][2].kernel_launches == 1


def test_exported_frame_keeps_its_size(worlds):
    """pt_frame_ipc_handle hands the BGRA frame to other processes: re-allocating it afterwards would leave them with a
    dangling mapping, so a render at another size is refused; a foreign pointer is refused as a resolve destination."""
    rc = render.RenderCore(0)
    rc.upload_scene(worlds(3).scene(64, 32))
    blob = rc.frame_ipc_handle(64, 32)
    assert len(blob) == render.IPC_HANDLE_BYTES and int.from_bytes(blob[72:76], "little") == 64 and int.from_bytes(blob[76:80], "little") == 32
    rc.render(64, 32, 2, seed=1)
    with pytest.raises(render.PtError) as e:
        rc.render(32, 32, 2, seed=1)
    assert e.value.code == -1 and "exported" in str(e.value)
    _, bdev = rc.device_buffers()
    with pytest.raises(render.PtError):
        rc.resolve_tiles_to(bdev.__cuda_array_interface__["data"][0], 0, 1)     # not a mapped peer frame
    rc.resolve_tiles_to(None, 0, 1)
    b, _, _ = rc.readback(True, False)
    rc.render(64, 32, 2, seed=1)
    assert np.array_equal(rc.readback(True, False)[0], b)
    rc.close()


def _p2p_worker(rank, world, port, out_dir):
    import os, sys
    import torch
    import torch.distributed as dist
    from conftest import ROOT
    sys.path.insert(0, ROOT)
    from pathtracer_b200 import dist as D
    from pathtracer_b200 import host
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    core = render.RenderCore(rank, stream=stream.cuda_stream)
    W, H, pp = 300, 170, 3
    core.upload_scene(host.HostWorld(5, scene_seed=7).scene(W, H))
    res = {}
    for mode in ("tiles", "tiles_p2p", "samples"):
        fr = D.ShardedFrame(core, W, H, mode)
        for seed in (4, 5):                              # two frames: the second exercises the frame-reuse ordering
            out = fr.render(pp, seed=seed)
            torch.cuda.synchronize()
            if rank == 0:
                res[f"{mode}_{seed}"] = out.cpu().numpy().view(np.uint32).copy()
        dist.barrier()
        fr.close()
    if rank == 0:
        np.savez(os.path.join(out_dir, "frames.npz"), **res)
    dist.barrier()
    core.close()
    dist.destroy_process_group()


def test_two_gpu_exchanges_give_the_one_gpu_frame(tmp_path, worlds):
    """2 ranks over NCCL on 2 GPUs: the gathered tiles, the tiles stored straight into rank 0's frame over NVLink peer
    memory (tiles_p2p) and the reduce-scatter of sample ranges all reproduce the 1-GPU frame (tiles: bit for bit)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    import socket
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mp.spawn(_p2p_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    got = np.load(tmp_path / "frames.npz")
    W, H, pp = 300, 170, 3
    rc = render.RenderCore(0)
    rc.upload_scene(worlds(5).scene(W, H))
    for seed in (4, 5):
        rc.render(W, H, pp, seed=seed)
        b, _, _ = rc.readback(True, False)
        assert np.array_equal(got[f"tiles_{seed}"], b)
        assert np.array_equal(got[f"tiles_p2p_{seed}"], b)
        assert (got[f"samples_{seed}"] == b).mean() > 0.999
    rc.close()


def test_largest_baseline_frame_properties(core, worlds):
    """C5's frame, 3840x2160 (the largest size BASELINE.json names): every pixel gets its samples, nothing is NaN or dropped,
    two of eight tile ranks reproduce their pixels bit for bit (the diagonal tile map at a width whose tile row, 120, is a
    multiple of 8), their shares are 1/8 of the frame, and no rank owns a vertical stripe."""
    from pathtracer_b200 import dist as D
    W, H, pp = 3840, 2160, 4
    core.upload_scene(worlds(5).scene(W, H))
    b, a, st = _frame(core, W, H, pp, seed=2)
    assert st.samples == W * H * pp * pp and (a[..., 3] == pp * pp).all() and np.isfinite(a).all()
    assert st.nan_dropped == 0 and (b >> 24 == 255).all()
    for r in (1, 6):
        _, ar, sr = _frame(core, W, H, pp, seed=2, tile_rank=r, tile_count=8)
        mine = ar[..., 3] > 0
        assert np.array_equal(ar[mine], a[mine]) and sr.samples == mine.sum() * pp * pp
        assert abs(mine.mean() - 1 / 8) < 0.005
        cols = np.unique(np.nonzero(mine.any(0))[0] // 32)
        assert len(cols) == 120                                   # tiles in every tile column: not a stripe
        tx, _ = D.tile_grid(W, H)
        ids = D.my_tiles(W, H, r, 8)
        t = ids[len(ids) // 2]
        assert mine[(t // tx) * 32:(t // tx) * 32 + 32, (t % tx) * 32:(t % tx) * 32 + 32].all()


@pytest.mark.parametrize("w", [1, 2, 3, 4, 5])
def test_no_sample_is_dropped_or_nan_on_any_world(core, worlds, w):
    """The reference redraws a NaN sample for ever (main.cpp:1068); the kernel redraws up to 8 times and then reports the
    sample in pt_stats.nan_dropped.  On the five worlds that counter stays 0 and every accumulator is finite."""
    W, H, pp = 320, 180, 8
    core.upload_scene(worlds(w).scene(W, H))
    _, a, st = _frame(core, W, H, pp, seed=17)
    assert st.nan_dropped == 0 and np.isfinite(a).all() and (a[..., 3] == pp * pp).all()
    assert st.nan_retries < 1e-4 * st.samples


def test_set_camera_keeps_the_scene(core, worlds):
    """pt_set_camera swaps the camera without touching the arena or the BVH: switching pinhole -> thin lens -> pinhole on an
    uploaded scene gives the images fresh uploads of those scenes give (the kernel variant follows the camera)."""
    W, H, pp = 120, 80, 4
    out = {}
    for lens in (False, True):
        core.upload_scene(worlds(5).scene(W, H, lens=lens))
        out[lens] = _frame(core, W, H, pp, seed=9)[1]
    core.upload_scene(worlds(5).scene(W, H, lens=False))
    for lens in (True, False, True):
        core.set_camera(worlds(5).scene(W, H, lens=lens).camera)
        assert np.array_equal(_frame(core, W, H, pp, seed=9)[1], out[lens])
