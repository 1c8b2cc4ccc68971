"""Synthetic scenes for the primitives the five worlds do not exercise (boxes above all): built directly as the
C-ABI's pt_scene from numpy arrays.  Keeps the arrays alive on the returned object."""
import ctypes as C

import numpy as np

from pathtracer_b200 import _ctypes_api as T
from pathtracer_b200 import host


class SynthScene:
    def __init__(self, seed=0, n_boxes=12, n_quads=3, n_spheres=4, plane=True, width=96, height=54, just_cosine=False):
        rng = np.random.default_rng(seed)
        n_mats = 6
        self.mats = np.zeros(n_mats, T.MATERIAL_DTYPE)
        m = self.mats
        m["alpha"], m["ior"] = 1.0, 1.45
        m["roughness"] = [1.0, 1.0, 0.6, 0.0, 0.3, 1.0]
        m["metalness"] = [0.0, 0.0, 0.0, 1.0, 0.7, 0.0]
        m["albedo"] = [[0, 0, 0], [0, 0, 0], [0.7, 0.3, 0.2], [0.9, 0.9, 0.9], [0.2, 0.6, 0.8], [0.6, 0.6, 0.6]]
        m["metal_color"] = [[0, 0, 0], [0, 0, 0], [1, 1, 1], [0.95, 0.8, 0.4], [0.8, 0.8, 0.9], [1, 1, 1]]
        m["emit"][0] = [0.3, 0.4, 0.6]          # material 0 = sky
        m["emit"][1] = [15, 15, 15]             # material 1 = the light
        self.spheres = np.zeros(1 + n_spheres, T.SPHERE_DTYPE)
        self.spheres[0] = ((0.0, 0.0, 6.0), 1.0, 1)          # spheres[0] is the sampled light
        for i in range(n_spheres):
            self.spheres[1 + i] = (tuple(rng.uniform(-3, 3, 2)) + (float(rng.uniform(0.3, 1.5)),), float(rng.uniform(0.2, 0.6)), int(rng.integers(2, n_mats)))
        self.planes = np.zeros(1 if plane else 0, T.PLANE_DTYPE)
        if plane:
            self.planes[0] = ((0.0, 0.0, 1.0), 0.0, 5)
        self.quads = np.zeros(n_quads, T.QUAD_DTYPE)
        for i in range(n_quads):
            p = rng.uniform(-3, 3, 3); p[2] = abs(p[2]) * 0.5 + 0.1
            u = rng.normal(size=3); v = np.cross(u, rng.normal(size=3))
            self.quads[i] = (tuple(p), tuple(u / np.linalg.norm(u) * rng.uniform(0.5, 1.5)), tuple(v / np.linalg.norm(v) * rng.uniform(0.5, 1.5)), int(rng.integers(2, n_mats)))
        self.box_origin = rng.uniform(-3, 3, (n_boxes, 3)).astype(np.float32)
        self.box_origin[:, 2] = np.abs(self.box_origin[:, 2]) * 0.6 + 0.3
        self.box_half = rng.uniform(0.1, 0.7, (n_boxes, 3)).astype(np.float32)
        self.aabbs = np.zeros(n_boxes, T.AABB_DTYPE)
        self.aabbs["origin"], self.aabbs["half_dim"] = self.box_origin, self.box_half
        self.aabbs["min"] = self.box_origin - self.box_half              # MakeAABB, ray_math.hpp:484-492
        self.aabbs["max"] = self.box_origin + self.box_half
        self.aabbs["mat"] = rng.integers(2, n_mats, n_boxes)
        s = self.scene = T.Scene()
        s.struct_size = C.sizeof(T.Scene)
        s.flags = T.SCENE_JUST_COSINE if just_cosine else 0
        s.n_materials, s.materials = n_mats, self.mats.ctypes.data
        s.n_spheres, s.spheres = len(self.spheres), self.spheres.ctypes.data
        s.n_planes, s.planes = len(self.planes), (self.planes.ctypes.data if plane else None)
        s.n_quads, s.quads = n_quads, (self.quads.ctypes.data if n_quads else None)
        s.n_aabbs, s.aabbs = n_boxes, (self.aabbs.ctypes.data if n_boxes else None)
        s.camera.use_pinhole = 1
        s.camera.pos, s.camera.target, s.camera.fov = T.V3(0, -9, 2.5), T.V3(0, 0, 0.8), 50.0
        s.camera.focal_distance, s.camera.aperture_radius = 5.0, 0.035
        host.define_camera(s.camera, width, height)
        self.width, self.height = width, height

    def random_rays(self, n, seed=1):
        rng = np.random.default_rng(seed)
        o = rng.uniform(-5, 5, (n, 3)).astype(np.float32)
        o[:, 2] = np.abs(o[:, 2]) * 0.7 + 0.05
        # a third of the rays start INSIDE a box (back faces, the early-exit paths of RayIntersectWithAABB2)
        k = rng.integers(0, max(1, len(self.box_origin)), n // 3)
        if len(self.box_origin):
            o[: n // 3] = self.box_origin[k] + rng.uniform(-0.9, 0.9, (n // 3, 3)).astype(np.float32) * self.box_half[k]
        d = rng.normal(size=(n, 3)).astype(np.float32)
        d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
        d[::17, 0] = 0.0                         # axis-parallel components: the degenerate slabs
        d[::23, 2] = 0.0
        d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
        return o, d


FACE_NORMALS = np.array([[0, 0, -1], [0, 0, 1], [-1, 0, 0], [1, 0, 0], [0, 1, 0], [0, -1, 0]], np.float32)   # main.cpp:540-543
