"""ctypes mirror of include/pt_api.h and include/pt_host.h (packed to 4 bytes, like the headers)."""
import ctypes as C

import numpy as np


class V3(C.Structure):
    _pack_ = 4
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float)]


class Material(C.Structure):
    _pack_ = 4
    _fields_ = [("alpha", C.c_float), ("ior", C.c_float), ("albedo_tex", C.c_int32), ("albedo", V3),
                ("metalness_tex", C.c_int32), ("metalness", C.c_float), ("metal_color", V3),
                ("roughness_tex", C.c_int32), ("roughness", C.c_float), ("normal_tex", C.c_int32), ("emit", V3)]


class Plane(C.Structure):
    _pack_ = 4
    _fields_ = [("n", V3), ("d", C.c_float), ("mat", C.c_uint32)]


class Quad(C.Structure):
    _pack_ = 4
    _fields_ = [("point", V3), ("u", V3), ("v", V3), ("mat", C.c_uint32)]


class Sphere(C.Structure):
    _pack_ = 4
    _fields_ = [("p", V3), ("r", C.c_float), ("mat", C.c_uint32)]


class Aabb(C.Structure):
    _pack_ = 4
    _fields_ = [("origin", V3), ("half_dim", V3), ("min", V3), ("max", V3), ("mat", C.c_uint32)]


class Camera(C.Structure):
    _pack_ = 4
    _fields_ = [("fov", C.c_float), ("focal_length", C.c_float), ("focal_distance", C.c_float),
                ("aperture_radius", C.c_float), ("use_pinhole", C.c_uint8), ("_pad", C.c_uint8 * 3),
                ("film_w", C.c_float), ("film_h", C.c_float), ("half_film_w", C.c_float), ("half_film_h", C.c_float),
                ("half_pixel_w", C.c_float), ("half_pixel_h", C.c_float),
                ("pos", V3), ("target", V3), ("frustum_center", V3), ("axis_z", V3), ("axis_x", V3), ("axis_y", V3)]


class Texture(C.Structure):
    _pack_ = 4
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("rgba8", C.c_void_p)]


class Scene(C.Structure):
    _pack_ = 4
    _fields_ = [("struct_size", C.c_uint32), ("flags", C.c_uint32),
                ("n_materials", C.c_uint32), ("n_spheres", C.c_uint32), ("n_planes", C.c_uint32),
                ("n_quads", C.c_uint32), ("n_aabbs", C.c_uint32), ("n_triangles", C.c_uint32),
                ("materials", C.c_void_p), ("spheres", C.c_void_p), ("planes", C.c_void_p),
                ("quads", C.c_void_p), ("aabbs", C.c_void_p), ("tri_points", C.c_void_p), ("tri_materials", C.c_void_p),
                ("textures", Texture * 4), ("camera", Camera)]


class Config(C.Structure):
    _pack_ = 4
    _fields_ = [("struct_size", C.c_uint32), ("device", C.c_int32), ("bvh_builder", C.c_int32),
                ("kernel", C.c_int32), ("stream", C.c_void_p)]


class RenderParams(C.Structure):
    _pack_ = 4
    _fields_ = [("struct_size", C.c_uint32), ("width", C.c_int32), ("height", C.c_int32), ("pp", C.c_int32),
                ("seed", C.c_uint64), ("sample_begin", C.c_uint32), ("sample_end", C.c_uint32),
                ("tile_rank", C.c_uint32), ("tile_count", C.c_uint32), ("flags", C.c_uint32)]


class Stats(C.Structure):
    _pack_ = 4
    _fields_ = [("samples", C.c_uint64), ("rays", C.c_uint64), ("nan_retries", C.c_uint64),
                ("render_ms", C.c_float), ("kernel_launches", C.c_uint32),
                ("bvh_nodes", C.c_uint32), ("bvh_prims", C.c_uint32),
                ("nan_dropped", C.c_uint64), ("texture_bytes_uploaded", C.c_uint64)]


class Tuning(C.Structure):
    _pack_ = 4
    _fields_ = [("struct_size", C.c_uint32), ("sample_lanes_log2", C.c_int32), ("shade_batch", C.c_int32),
                ("coop_max", C.c_int32), ("tri_leaf", C.c_int32), ("sph_leaf", C.c_int32), ("sah_sweep", C.c_int32),
                ("wf_pixels_log2", C.c_int32)]


class HostOptions(C.Structure):
    _fields_ = [("threads", C.c_int32), ("pp", C.c_int32), ("world", C.c_int32), ("lens", C.c_int32),
                ("normals", C.c_int32), ("metalness", C.c_int32), ("roughness", C.c_int32), ("help", C.c_int32),
                ("width", C.c_int32), ("height", C.c_int32), ("gpus", C.c_int32), ("shard_samples", C.c_int32),
                ("seed", C.c_uint64), ("scene_seed", C.c_uint32), ("lbvh", C.c_int32), ("kernel", C.c_int32),
                ("out_path", C.c_char * 512), ("res_dir", C.c_char * 512), ("system_cores", C.c_int32),
                ("snapshots", C.c_int32), ("hdr_dump", C.c_int32)]


HIT_DTYPE = np.dtype([("t", "<f4"), ("kind", "<i4"), ("index", "<i4"), ("mat", "<u4"), ("normal", "<f4", 3)])

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

assert C.sizeof(Material) == 68 and MATERIAL_DTYPE.itemsize == 68
assert C.sizeof(Plane) == 20 and C.sizeof(Quad) == 40 and C.sizeof(Sphere) == 20 and C.sizeof(Aabb) == 52
assert C.sizeof(Camera) == 116 and CAMERA_DTYPE.itemsize == 116
assert HIT_DTYPE.itemsize == 28

SCENE_NORMAL_MAPS, SCENE_METALNESS_MAPS, SCENE_ROUGHNESS_MAPS, SCENE_JUST_COSINE = 1, 2, 4, 8
SCENE_ALL_MAPS = 7
SCENE_KEEP_TEXTURES = 1 << 8
RENDER_CLEAR = 1
BVH_HOST_SAH, BVH_DEVICE_LBVH = 0, 1
KERNEL_AUTO, KERNEL_MEGA, KERNEL_WAVEFRONT = 0, 1, 2
HIT_MISS, HIT_SPHERE, HIT_QUAD, HIT_PLANE, HIT_TRIANGLE, HIT_AABB = range(6)
