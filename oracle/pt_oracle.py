"""ctypes view of oracle/libpt_oracle.so, the plain-C CPU restatement of the reference hot path.

TEST INFRASTRUCTURE ONLY: the checker for tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product (pathtracer_b200.render) never imports this.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libpt_oracle.so")

RNG_MT19937, RNG_PHILOX = 0, 1

HIT_DTYPE = np.dtype([("t", "<f4"), ("kind", "<i4"), ("index", "<i4"), ("mat", "<u4"), ("normal", "<f4", 3)])

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is not built (make -C oracle)")
        L = C.CDLL(LIB_PATH)
        L.pto_philox_uniform.restype = C.c_float
        L.pto_philox_uniform.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
        L.pto_intersect.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
        L.pto_raycast.argtypes = [C.c_void_p, C.c_int, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
        L.pto_render.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64,
                                 C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_uint32,
                                 C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.pto_render_threads.restype = C.c_double
        L.pto_render_threads.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p]
        L.pto_resolve.argtypes = [C.c_uint32, C.c_void_p, C.c_void_p]
        L.pto_sample_texture.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_void_p]
        L.pto_primary_rays.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.pto_seed_mt.argtypes = [C.c_uint32]
        L.pto_random_mt.argtypes = [C.c_uint32, C.c_void_p]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def seed_mt(s):
    lib().pto_seed_mt(s)


def random_mt(n):
    out = np.zeros(n, np.float32)
    lib().pto_random_mt(n, _p(out))
    return out


def philox_uniform(seed, a, b, tag, k):
    return lib().pto_philox_uniform(seed, a, b, tag, k)


def intersect(scene, o, d):
    o = np.ascontiguousarray(o, np.float32).reshape(-1, 3)
    d = np.ascontiguousarray(d, np.float32).reshape(-1, 3)
    hits = np.zeros(o.shape[0], HIT_DTYPE)
    lib().pto_intersect(C.addressof(scene), o.shape[0], _p(o), _p(d), _p(hits))
    return hits


def raycast(scene, o, d, rng_mode=RNG_PHILOX, seed=1):
    o = np.ascontiguousarray(o, np.float32).reshape(-1, 3)
    d = np.ascontiguousarray(d, np.float32).reshape(-1, 3)
    rgb = np.zeros((o.shape[0], 3), np.float32)
    lib().pto_raycast(C.addressof(scene), rng_mode, seed, o.shape[0], _p(o), _p(d), _p(rgb))
    return rgb


def render(scene, width, height, pp, rng_mode=RNG_PHILOX, seed=1, window=None, s_begin=0, s_end=0):
    x0, y0, w, h = window if window else (0, 0, width, height)
    s = np.zeros((h, w, 3), np.float64)
    s2 = np.zeros((h, w, 3), np.float64)
    mf = np.zeros((h, w, 3), np.float32)
    ldr = np.zeros((h, w), np.uint32)
    lib().pto_render(C.addressof(scene), width, height, pp, rng_mode, seed, x0, y0, w, h, s_begin, s_end,
                     _p(s), _p(s2), _p(mf), _p(ldr))
    n = (s_end if s_end else pp * pp) - s_begin
    return dict(sum=s, sumsq=s2, mean_f=mf, ldr=ldr, spp=n)


def render_threads(scene, width, height, pp, seed=1, threads=1):
    mf = np.zeros((height, width, 3), np.float32)
    ldr = np.zeros((height, width), np.uint32)
    sec = lib().pto_render_threads(C.addressof(scene), width, height, pp, seed, threads, _p(mf), _p(ldr))
    return sec, mf, ldr


def resolve(rgb):
    rgb = np.ascontiguousarray(rgb, np.float32).reshape(-1, 3)
    out = np.zeros(rgb.shape[0], np.uint32)
    lib().pto_resolve(rgb.shape[0], _p(rgb), _p(out))
    return out


def sample_texture(scene, idx, u, v):
    out = np.zeros(3, np.float32)
    lib().pto_sample_texture(C.addressof(scene), idx, u, v, _p(out))
    return out


def primary_rays(scene, width, height):
    o = np.zeros((width * height, 3), np.float32)
    d = np.zeros((width * height, 3), np.float32)
    lib().pto_primary_rays(C.addressof(scene), width, height, _p(o), _p(d))
    return o, d
