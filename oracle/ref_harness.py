"""ctypes view of oracle/_ref/libref_harness.so -- the reference renderer itself (its own sources
compiled where they lie under /root/reference by `make -C oracle ref`).

TEST INFRASTRUCTURE ONLY: imported by tests/, tools/ that generate golden fixtures,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.  Never by the product.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_ref", "libref_harness.so")
RUN_DIR = os.path.join(HERE, "_ref", "run")
EXE_PATH = os.path.join(HERE, "_ref", "pathtracer_ref")

FLAG_LENS, FLAG_NO_NORMALS, FLAG_NO_METALNESS, FLAG_NO_ROUGHNESS = 1, 2, 4, 8

MATERIAL_DTYPE = np.dtype([
    ("alpha", "<f4"), ("ior", "<f4"), ("albedo_tex", "<i4"), ("albedo", "<f4", 3),
    ("metalness_tex", "<i4"), ("metalness", "<f4"), ("metal_color", "<f4", 3),
    ("roughness_tex", "<i4"), ("roughness", "<f4"), ("normal_tex", "<i4"), ("emit", "<f4", 3)])
SPHERE_DTYPE = np.dtype([("p", "<f4", 3), ("r", "<f4"), ("mat", "<u4")])
PLANE_DTYPE = np.dtype([("n", "<f4", 3), ("d", "<f4"), ("mat", "<u4")])
QUAD_DTYPE = np.dtype([("point", "<f4", 3), ("u", "<f4", 3), ("v", "<f4", 3), ("mat", "<u4")])
AABB_DTYPE = np.dtype([("origin", "<f4", 3), ("half_dim", "<f4", 3), ("min", "<f4", 3), ("max", "<f4", 3), ("mat", "<u4")])
CAMERA_DTYPE = np.dtype([
    ("fov", "<f4"), ("focal_length", "<f4"), ("focal_distance", "<f4"), ("aperture_radius", "<f4"),
    ("use_pinhole", "u1"), ("_pad", "u1", 3),
    ("film_w", "<f4"), ("film_h", "<f4"), ("half_film_w", "<f4"), ("half_film_h", "<f4"),
    ("half_pixel_w", "<f4"), ("half_pixel_h", "<f4"),
    ("pos", "<f4", 3), ("target", "<f4", 3), ("frustum_center", "<f4", 3),
    ("axis_z", "<f4", 3), ("axis_x", "<f4", 3), ("axis_y", "<f4", 3)])


def available():
    return os.path.exists(LIB_PATH) and os.path.isdir(RUN_DIR)


def _fp(a):
    return a.ctypes.data_as(C.c_void_p)


class RefHarness:
    """One process-wide instance (the reference keeps its scene in file-static globals)."""
    _lib = None

    def __init__(self):
        if RefHarness._lib is None:
            if not available():
                raise RuntimeError("oracle/_ref is not built (make -C oracle ref; needs /root/reference)")
            lib = C.CDLL(LIB_PATH)
            lib.ref_load.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint]
            lib.ref_load.restype = C.c_int
            lib.ref_render_threads.restype = C.c_double
            lib.ref_render_threads.argtypes = [C.c_int, C.c_void_p]
            lib.ref_resolve.restype = C.c_uint
            lib.ref_resolve.argtypes = [C.c_float, C.c_float, C.c_float]
            lib.ref_sample_texture.argtypes = [C.c_int, C.c_float, C.c_float, C.c_void_p]
            RefHarness._lib = lib
            assert lib.ref_sizeof(0) == MATERIAL_DTYPE.itemsize
            assert lib.ref_sizeof(1) == SPHERE_DTYPE.itemsize
            assert lib.ref_sizeof(2) == PLANE_DTYPE.itemsize
            assert lib.ref_sizeof(3) == QUAD_DTYPE.itemsize
            assert lib.ref_sizeof(4) == AABB_DTYPE.itemsize
            assert lib.ref_sizeof(5) == CAMERA_DTYPE.itemsize
        self.lib = RefHarness._lib
        self.W = self.H = self.pp = 0

    def load(self, world, W, H, pp=4, flags=0, scene_seed=7):
        cwd = os.getcwd()
        try:
            rc = self.lib.ref_load(RUN_DIR.encode(), world, W, H, pp, flags, scene_seed)
        finally:
            os.chdir(cwd)
        if rc != 0:
            raise RuntimeError(f"ref_load failed: {rc}")
        self.W, self.H, self.pp, self.world = W, H, pp, world
        return self

    def seed(self, s):
        self.lib.ref_seed(C.c_uint(s))

    def counts(self):
        c = (C.c_int * 8)()
        self.lib.ref_get_counts(c)
        return list(c)

    def scene(self):
        c = self.counts()
        mats = np.zeros(c[0], MATERIAL_DTYPE)
        sph = np.zeros(c[1], SPHERE_DTYPE)
        pla = np.zeros(c[2], PLANE_DTYPE)
        qua = np.zeros(c[3], QUAD_DTYPE)
        aab = np.zeros(c[4], AABB_DTYPE)
        pts = np.zeros((c[5], 3), np.float32)
        pmat = np.zeros(c[5], np.int32)
        cam = np.zeros(1, CAMERA_DTYPE)
        self.lib.ref_get_scene(_fp(mats), _fp(sph), _fp(pla), _fp(qua), _fp(aab), _fp(pts), _fp(pmat), _fp(cam))
        return dict(materials=mats, spheres=sph, planes=pla, quads=qua, aabbs=aab,
                    tri_points=pts.reshape(-1, 9), tri_materials=pmat[0::3].copy(), point_materials=pmat,
                    camera=cam[0])

    def texture(self, idx):
        w, h, p = C.c_int(), C.c_int(), C.POINTER(C.c_float)()
        if self.lib.ref_get_texture(idx, C.byref(w), C.byref(h), C.byref(p)) != 0:
            return None
        return np.ctypeslib.as_array(p, shape=(h.value, w.value, 3)).copy()

    def intersect(self, o, d, with_ids=True):
        o = np.ascontiguousarray(o, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(d, np.float32).reshape(-1, 3)
        n = o.shape[0]
        t = np.zeros(n, np.float32)
        mat = np.zeros(n, np.uint32)
        nor = np.zeros((n, 3), np.float32)
        kind = np.zeros(n, np.int32)
        idx = np.zeros(n, np.int32)
        self.lib.ref_intersect(n, _fp(o), _fp(d), _fp(t), _fp(mat), _fp(nor),
                               _fp(kind) if with_ids else None, _fp(idx) if with_ids else None)
        return dict(t=t, mat=mat, normal=nor, kind=kind, index=idx)

    def raycast(self, o, d):
        o = np.ascontiguousarray(o, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(d, np.float32).reshape(-1, 3)
        rgb = np.zeros((o.shape[0], 3), np.float32)
        self.lib.ref_raycast(o.shape[0], _fp(o), _fp(d), _fp(rgb))
        return rgb

    def primary_rays(self):
        o = np.zeros((self.H * self.W, 3), np.float32)
        d = np.zeros((self.H * self.W, 3), np.float32)
        self.lib.ref_primary_rays(_fp(o), _fp(d))
        return o, d

    def render_texel(self, x0, y0, w, h):
        out = np.zeros((h, w), np.uint32)
        self.lib.ref_render_texel(x0, y0, w, h, _fp(out))
        return out

    def render_hdr(self, x0=0, y0=0, w=None, h=None):
        w = self.W if w is None else w
        h = self.H if h is None else h
        s = np.zeros((h, w, 3), np.float64)
        s2 = np.zeros((h, w, 3), np.float64)
        mf = np.zeros((h, w, 3), np.float32)
        ldr = np.zeros((h, w), np.uint32)
        self.lib.ref_render_hdr(x0, y0, w, h, _fp(s), _fp(s2), _fp(mf), _fp(ldr))
        return dict(sum=s, sumsq=s2, mean_f=mf, ldr=ldr, spp=self.pp * self.pp)

    def render_threads(self, threads):
        px = np.zeros((self.H, self.W), np.uint32)
        cwd = os.getcwd()
        os.chdir(RUN_DIR)
        try:
            sec = self.lib.ref_render_threads(threads, _fp(px))
        finally:
            os.chdir(cwd)
        return sec, px

    def resolve(self, rgb):
        rgb = np.asarray(rgb, np.float32).reshape(-1, 3)
        return np.array([self.lib.ref_resolve(float(r), float(g), float(b)) for r, g, b in rgb], np.uint32)

    def random(self, n):
        out = np.zeros(n, np.float32)
        self.lib.ref_random(n, _fp(out))
        return out

    def aabb_test(self, o, d, origins, half_dims):
        """RayIntersectWithAABB2 over boxes MakeAABB(origin, half_dim), nearest accepted as main.cpp:529-553."""
        o = np.ascontiguousarray(o, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(d, np.float32).reshape(-1, 3)
        bo = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        bh = np.ascontiguousarray(half_dims, np.float32).reshape(-1, 3)
        t = np.zeros(o.shape[0], np.float32)
        idx = np.zeros(o.shape[0], np.int32)
        face = np.zeros(o.shape[0], np.int32)
        self.lib.ref_aabb_test(o.shape[0], _fp(o), _fp(d), bo.shape[0], _fp(bo), _fp(bh), _fp(t), _fp(idx), _fp(face))
        return dict(t=t, index=idx, face=face)

    def sample_texture(self, idx, u, v):
        out = np.zeros(3, np.float32)
        self.lib.ref_sample_texture(idx, C.c_float(u), C.c_float(v), _fp(out))
        return out
