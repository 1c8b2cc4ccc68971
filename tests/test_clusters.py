"""CPU checks of the cluster cut behind the warp-cooperative closest hit (pathtracer_b200/csrc/pt_cluster.h):
the clusters tile the leaf order, their boxes hold their primitives, and a lane-by-lane emulation of
pt_coop_trace returns the same (t, id) as the BVH2 walk and as brute force -- the argument for why the
PT_F_COOP kernel variants stay bit-identical to the default ones (tests/test_gpu_properties.py checks that on the GPU)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT
from pathtracer_b200 import host


class Stats(C.Structure):
    _fields_ = [("n_clusters", C.c_uint32), ("max_count", C.c_uint32), ("min_count", C.c_uint32), ("partition_ok", C.c_uint32),
                ("contain_ok", C.c_uint32), ("n_rays", C.c_uint32), ("mismatch_walk", C.c_uint32), ("mismatch_coop", C.c_uint32),
                ("n_hits", C.c_uint32), ("walk_steps_per_ray", C.c_double), ("coop_rounds_per_ray", C.c_double),
                ("coop_clusters_per_ray", C.c_double), ("firsts", C.c_float * 32), ("counts", C.c_float * 32)]


@pytest.fixture(scope="module")
def checker(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("cluster") / "libcluster_check.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-I" + os.path.join(ROOT, "pathtracer_b200", "csrc"),
                           "-I" + os.path.join(ROOT, "include"), "-o", so, os.path.join(ROOT, "tests", "cpp", "cluster_check.cpp")])
    L = C.CDLL(so)
    L.cluster_check.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_uint32, C.c_uint32, C.c_float, C.POINTER(Stats)]

    def run(boxes, tris, max_leaf, n_rays=0, seed=1, spread=1.0):
        boxes = np.ascontiguousarray(boxes, np.float32)
        tp = None
        if tris is not None:
            tris = np.ascontiguousarray(tris, np.float32)
            tp = tris.ctypes.data_as(C.c_void_p)
        st = Stats()
        rc = L.cluster_check(boxes.ctypes.data_as(C.c_void_p), tp, len(boxes), max_leaf, n_rays, seed, spread, C.byref(st))
        return rc, st
    return run


def tri_boxes(tris):
    p = tris.reshape(-1, 3, 3)
    return np.concatenate([p.min(1), p.max(1)], 1)


def test_mario_mesh_cut_and_cooperative_trace(checker, worlds):
    tris = host.scene_arrays(worlds(5).scene(64, 36))["tri_points"].copy()
    assert len(tris) == 752
    for spread, min_hit_frac in [(1.0, 0.05), (2.5, 0.0)]:          # rays into the box; rays that mostly cross it and miss
        rc, st = checker(tri_boxes(tris), tris, 4, n_rays=20000, seed=3, spread=spread)
        assert rc == 0 and st.partition_ok and st.contain_ok
        assert 8 <= st.n_clusters <= 32 and st.max_count <= 128
        assert st.n_rays == 20000 and st.mismatch_walk == 0 and st.mismatch_coop == 0
        assert st.n_hits >= min_hit_frac * st.n_rays
        # the point of the cut: a handful of 32-wide rounds instead of tens of node steps
        assert st.coop_rounds_per_ray < 0.5 * st.walk_steps_per_ray


@pytest.mark.parametrize("n,max_leaf,seed", [(1, 4, 1), (3, 4, 2), (33, 4, 3), (500, 4, 4), (1900, 4, 5), (300, 1, 6), (2000, 8, 7)])
def test_random_triangle_soups(checker, n, max_leaf, seed):
    rng = np.random.default_rng(seed)
    ctr = rng.uniform(-1, 1, (n, 1, 3)).astype(np.float32)
    tris = (ctr + rng.normal(0, 0.08, (n, 3, 3)).astype(np.float32)).reshape(n, 9)
    rc, st = checker(tri_boxes(tris), tris, max_leaf, n_rays=4000, seed=seed, spread=1.5)
    assert rc == 0, "meshes up to ~2000 primitives have a cut (two sibling clusters always hold > 128 primitives together)"
    assert st.partition_ok and st.contain_ok and st.n_clusters <= 32 and st.max_count <= 128
    assert st.mismatch_walk == 0 and st.mismatch_coop == 0


def test_too_large_a_mesh_has_no_single_level_cut(checker):
    rng = np.random.default_rng(9)
    n = 6000                                                       # > 32 clusters x 128 primitives
    tris = (rng.uniform(-1, 1, (n, 1, 3)) + rng.normal(0, 0.02, (n, 3, 3))).astype(np.float32).reshape(n, 9)
    rc, st = checker(tri_boxes(tris), None, 4)
    assert rc == 1 and st.n_clusters == 0                          # pt_render then keeps the default kernel variant


@pytest.mark.parametrize("world", [2, 4])
def test_sphere_worlds_have_a_cut(checker, worlds, world):
    sph = host.scene_arrays(worlds(world).scene(64, 36))["spheres"]
    c = sph["p"].astype(np.float32)
    r = sph["r"].astype(np.float32)[:, None]
    small = (r[:, 0] <= 16 * np.median(r))                          # the library keeps the huge spheres in its linear list
    boxes = np.concatenate([c - r, c + r], 1)[small]
    rc, st = checker(boxes, None, 2)
    assert rc == 0 and st.partition_ok and st.contain_ok and 2 <= st.n_clusters <= 32


def test_fast_sah_builder_builds_the_identical_tree(tmp_path_factory):
    """pt_bvh.h sorts with a stable radix sort, skips the axis the parent already sorted and sweeps SSE boxes (the BVH build is
    most of the host time of pt_upload_scene).  Against the plain formulation of the same algorithm (tests/cpp/bvh_check.cpp:
    std::stable_sort per axis per node, scalar boxes) the tree must be IDENTICAL -- nodes bit for bit, primitive order, depth --
    on random boxes, on the worlds' own primitives, and on inputs full of ties (grids, coincident centroids, -0.0)."""
    so = str(tmp_path_factory.mktemp("bvh") / "libbvh_check.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-I" + os.path.join(ROOT, "pathtracer_b200", "csrc"),
                           "-I" + os.path.join(ROOT, "include"), "-o", so, os.path.join(ROOT, "tests", "cpp", "bvh_check.cpp")])
    L = C.CDLL(so)
    L.bvh_check.argtypes = [C.c_void_p, C.c_uint32, C.c_int]

    def check(boxes, leaf):
        boxes = np.ascontiguousarray(boxes, np.float32)
        assert L.bvh_check(boxes.ctypes.data_as(C.c_void_p), len(boxes), leaf) == 0, (len(boxes), leaf)

    rng = np.random.default_rng(3)
    for n in (5, 47, 48, 49, 200, 752, 3000):
        c = rng.uniform(-3, 3, (n, 3))
        h = rng.uniform(0.01, 0.3, (n, 3))
        for leaf in (1, 3, 4):
            check(np.hstack([c - h, c + h]), leaf)
    # ties everywhere: a grid of equal spheres in a plane (world 2 / 4 style), coincident boxes, zero and negative-zero centroids
    g = np.stack(np.meshgrid(np.arange(11) * 0.5, np.arange(11) * 0.5, [0.2]), -1).reshape(-1, 3)
    check(np.hstack([g - 0.2, g + 0.2]), 3)
    same = np.tile([[-1, -1, -1, 1, 1, 1]], (70, 1)).astype(np.float32)
    check(same, 2)
    z = np.zeros((100, 6), np.float32)
    z[:, 0] = -rng.integers(0, 3, 100); z[:, 3] = -z[:, 0]                   # centroids exactly +0 ... and
    z[::2, 1], z[::2, 4] = -0.0, -0.0                                        # ... -0.0 on another axis
    z[:, 2] = rng.integers(0, 4, 100); z[:, 5] = z[:, 2] + 1
    check(z, 2)
    # the worlds' own trees: Mario's triangles, the sphere fields of worlds 2 and 4
    s5 = host.scene_arrays(host.HostWorld(5).scene(64, 36))
    t = s5["tri_points"].reshape(-1, 3, 3)
    check(np.hstack([t.min(1), t.max(1)]), 4)
    for w in (2, 4):
        sp = host.scene_arrays(host.HostWorld(w).scene(64, 36))["spheres"]
        c, r = sp["p"], sp["r"][:, None]
        keep = sp["r"] < 16 * np.median(sp["r"])
        check(np.hstack([c - r, c + r])[keep], 3)
