"""Python view of libpt_b200.so, the CUDA render core behind include/pt_api.h.

This is the product path: pt_upload_scene / pt_render / pt_readback (+ the parity exports).  It
never imports anything from oracle/ and has no CPU fallback: a missing library or a missing CUDA
device raises.
"""
import ctypes as C
import os

import numpy as np

from . import _ctypes_api as T

_HERE = os.path.dirname(os.path.abspath(__file__))
# PT_B200_LIB: another build of the same library (kernel experiments: tools/build_variants.sh)
LIB_PATH = os.environ.get("PT_B200_LIB") or os.path.join(_HERE, "libpt_b200.so")

NOALLOC_MAX_COLORS = 65536  # PT_NOALLOC_MAX_COLORS
IPC_HANDLE_BYTES = 80       # PT_IPC_HANDLE_BYTES
_lib = None


class PtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"pt error {code}: {msg}")
        self.code = code


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is not built (python -c 'import __graft_entry__ as g; g.build()'); "
                               "the render core has no CPU fallback")
        L = C.CDLL(LIB_PATH)
        L.pt_last_error.restype = C.c_char_p
        L.pt_last_error.argtypes = [C.c_void_p]
        L.pt_create.argtypes = [C.POINTER(T.Config), C.POINTER(C.c_void_p)]
        L.pt_destroy.argtypes = [C.c_void_p]
        L.pt_upload_scene.argtypes = [C.c_void_p, C.POINTER(T.Scene)]
        L.pt_set_camera.argtypes = [C.c_void_p, C.POINTER(T.Camera)]
        L.pt_set_tuning.argtypes = [C.c_void_p, C.POINTER(T.Tuning)]
        L.pt_resolve_pack_tiles.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint32)]
        L.pt_unpack_gathered.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]
        L.pt_resolve_range.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]
        L.pt_mark_resolved.argtypes = [C.c_void_p]
        L.pt_tile_owner.argtypes = [C.c_int32, C.c_int32, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint32)]
        L.pt_tile_owner.restype = C.c_uint32
        L.pt_tile_slots.argtypes = [C.c_int32, C.c_int32, C.c_uint32]
        L.pt_tile_slots.restype = C.c_uint32
        L.pt_render.argtypes = [C.c_void_p, C.POINTER(T.RenderParams)]
        L.pt_synchronize.argtypes = [C.c_void_p]
        L.pt_resolve.argtypes = [C.c_void_p]
        L.pt_readback.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(T.Stats)]
        L.pt_device_buffers.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                        C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        L.pt_set_stream.argtypes = [C.c_void_p, C.c_void_p]
        L.pt_flush_l2.argtypes = [C.c_void_p]
        L.pt_gather_tiles_from.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]
        L.pt_add_accum_from.argtypes = [C.c_void_p, C.c_void_p]
        L.pt_pack_tiles.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint32)]
        L.pt_unpack_tiles.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]
        L.pt_frame_ipc_handle.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]
        L.pt_open_peer_frame.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p)]
        L.pt_close_peer_frame.argtypes = [C.c_void_p, C.c_void_p]
        L.pt_resolve_tiles_to.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]
        L.pt_trace_rays.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
        L.pt_primary_hits.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]
        L.pt_trace_paths.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
        L.pt_resolve_colors.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]
        L.pt_resolve_colors_noalloc.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]
        L.pt_measure_fp32_peak.argtypes = [C.c_void_p, C.POINTER(C.c_float)]
        _lib = L
    return _lib


EXPORTS = ["pt_api_version", "pt_create", "pt_destroy", "pt_last_error", "pt_upload_scene", "pt_set_camera",
           "pt_render", "pt_synchronize", "pt_resolve", "pt_readback", "pt_device_buffers", "pt_set_stream",
           "pt_pack_tiles", "pt_unpack_tiles", "pt_gather_tiles_from", "pt_add_accum_from", "pt_trace_rays", "pt_primary_hits", "pt_trace_paths",
           "pt_resolve_colors", "pt_resolve_colors_noalloc", "pt_measure_fp32_peak", "pt_flush_l2",
           "pt_frame_ipc_handle", "pt_open_peer_frame", "pt_close_peer_frame", "pt_resolve_tiles_to",
           "pt_tile_owner", "pt_tile_slots", "pt_set_tuning", "pt_resolve_pack_tiles", "pt_unpack_gathered", "pt_resolve_range", "pt_mark_resolved"]


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class DeviceArray:
    """Zero-copy handle on a device buffer owned by the context (for torch.as_tensor / NCCL)."""

    def __init__(self, ptr, shape, typestr, owner):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 3, "strides": None}
        self._owner = owner


class RenderCore:
    """One context on one GPU.  Not re-entrant; use from one host thread at a time."""

    def __init__(self, device=0, bvh=T.BVH_HOST_SAH, kernel=T.KERNEL_AUTO, stream=None):
        self._h = C.c_void_p()
        cfg = T.Config(C.sizeof(T.Config), device, bvh, kernel, stream)
        rc = lib().pt_create(C.byref(cfg), C.byref(self._h))
        if rc != 0:
            raise PtError(rc, lib().pt_last_error(None).decode())
        self.width = self.height = 0

    def _ck(self, rc):
        if rc != 0:
            raise PtError(rc, lib().pt_last_error(self._h).decode())

    def close(self):
        if self._h:
            lib().pt_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload_scene(self, scene):
        self._ck(lib().pt_upload_scene(self._h, C.byref(scene)))

    def set_camera(self, cam):
        self._ck(lib().pt_set_camera(self._h, C.byref(cam)))

    def set_tuning(self, sample_lanes_log2=-1, shade_batch=0, coop_max=-1, tri_leaf=0, sph_leaf=0, sah_sweep=-1,
                   wf_pixels_log2=0):
        """Scheduling knobs (pt_tuning): never change a sample.  coop_max / leaf sizes apply to the NEXT upload."""
        t = T.Tuning(C.sizeof(T.Tuning), sample_lanes_log2, shade_batch, coop_max, tri_leaf, sph_leaf, sah_sweep,
                     wf_pixels_log2)
        self._ck(lib().pt_set_tuning(self._h, C.byref(t)))

    def set_stream(self, cuda_stream_ptr):
        self._ck(lib().pt_set_stream(self._h, cuda_stream_ptr))

    def render(self, width, height, pp, seed=1, sample_begin=0, sample_end=0, tile_rank=0, tile_count=1, clear=True):
        p = T.RenderParams(C.sizeof(T.RenderParams), width, height, pp, seed, sample_begin, sample_end,
                           tile_rank, tile_count, T.RENDER_CLEAR if clear else 0)
        self._ck(lib().pt_render(self._h, C.byref(p)))
        self.width, self.height = width, height

    def synchronize(self):
        self._ck(lib().pt_synchronize(self._h))

    def resolve(self):
        self._ck(lib().pt_resolve(self._h))

    def readback(self, bgra=True, accum=False, out_bgra=None, out_accum=None):
        """Returns (bgra [H,W] uint32 or None, accum [H,W,4] float32 or None, Stats)."""
        b = a = None
        if bgra:
            b = out_bgra if out_bgra is not None else np.empty((self.height, self.width), np.uint32)
        if accum:
            a = out_accum if out_accum is not None else np.empty((self.height, self.width, 4), np.float32)
        st = T.Stats()
        self._ck(lib().pt_readback(self._h, _p(b), _p(a), C.byref(st)))
        return b, a, st

    def device_buffers(self):
        pa, pb, w, h = C.c_void_p(), C.c_void_p(), C.c_int32(), C.c_int32()
        self._ck(lib().pt_device_buffers(self._h, C.byref(pa), C.byref(pb), C.byref(w), C.byref(h)))
        return (DeviceArray(pa.value, (h.value, w.value, 4), "<f4", self),
                DeviceArray(pb.value, (h.value, w.value), "<i4", self))   # BGRA words viewed as int32 for torch

    def pack_tiles(self, dst_ptr, tile_rank, tile_count):
        n = C.c_uint32()
        self._ck(lib().pt_pack_tiles(self._h, dst_ptr, tile_rank, tile_count, C.byref(n)))
        return n.value

    def unpack_tiles(self, src_ptr, tile_rank, tile_count):
        self._ck(lib().pt_unpack_tiles(self._h, src_ptr, tile_rank, tile_count))

    def resolve_pack_tiles(self, dst_ptr, tile_rank, tile_count):
        n = C.c_uint32()
        self._ck(lib().pt_resolve_pack_tiles(self._h, dst_ptr, tile_rank, tile_count, C.byref(n)))
        return n.value

    def unpack_gathered(self, src_ptr, tile_count, stride_words):
        self._ck(lib().pt_unpack_gathered(self._h, src_ptr, tile_count, stride_words))

    def resolve_range(self, accum_ptr, bgra_ptr, n_pixels):
        self._ck(lib().pt_resolve_range(self._h, accum_ptr, bgra_ptr, n_pixels))

    def mark_resolved(self):
        self._ck(lib().pt_mark_resolved(self._h))

    def flush_l2(self):
        self._ck(lib().pt_flush_l2(self._h))

    def gather_tiles_from(self, src, tile_rank, tile_count):
        """Single-process multi-GPU: pulls `src`'s tiles into this context's frame by a peer copy."""
        self._ck(lib().pt_gather_tiles_from(self._h, src._h, tile_rank, tile_count))

    def add_accum_from(self, src):
        """Single-process multi-GPU: accum += src.accum (sample-range sharding)."""
        self._ck(lib().pt_add_accum_from(self._h, src._h))

    # ---- direct tile exchange over peer memory (one process per GPU) ----
    def frame_ipc_handle(self, width, height):
        """PT_IPC_HANDLE_BYTES blob naming this context's BGRA frame (allocated here if needed)."""
        buf = (C.c_uint8 * IPC_HANDLE_BYTES)()
        self._ck(lib().pt_frame_ipc_handle(self._h, width, height, buf))
        self.width, self.height = width, height
        return bytes(buf)

    def open_peer_frame(self, handle):
        p = C.c_void_p()
        buf = (C.c_uint8 * IPC_HANDLE_BYTES).from_buffer_copy(handle)
        self._ck(lib().pt_open_peer_frame(self._h, buf, C.byref(p)))
        return p.value

    def close_peer_frame(self, remote):
        self._ck(lib().pt_close_peer_frame(self._h, remote))

    def resolve_tiles_to(self, dst_ptr, tile_rank, tile_count):
        """Resolve this rank's tiles into dst_ptr (a mapped peer frame) or, with None, into the own frame."""
        self._ck(lib().pt_resolve_tiles_to(self._h, dst_ptr, tile_rank, tile_count))

    # ---- parity / debug exports ----
    def trace_rays(self, o, d):
        o = np.ascontiguousarray(o, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(d, np.float32).reshape(-1, 3)
        hits = np.zeros(o.shape[0], T.HIT_DTYPE)
        self._ck(lib().pt_trace_rays(self._h, o.shape[0], _p(o), _p(d), _p(hits)))
        return hits

    def primary_hits(self, width, height):
        hits = np.zeros(width * height, T.HIT_DTYPE)
        self._ck(lib().pt_primary_hits(self._h, width, height, _p(hits)))
        return hits

    def trace_paths(self, o, d, seed=1):
        o = np.ascontiguousarray(o, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(d, np.float32).reshape(-1, 3)
        rgb = np.zeros((o.shape[0], 3), np.float32)
        self._ck(lib().pt_trace_paths(self._h, o.shape[0], _p(o), _p(d), seed, _p(rgb)))
        return rgb

    def resolve_colors(self, rgb):
        rgb = np.ascontiguousarray(rgb, np.float32).reshape(-1, 3)
        out = np.zeros(rgb.shape[0], np.uint32)
        self._ck(lib().pt_resolve_colors(self._h, rgb.shape[0], _p(rgb), _p(out)))
        return out


    def resolve_colors_noalloc(self, rgb):
        """pt_resolve_colors_noalloc: the same result with no device allocation (n <= NOALLOC_MAX_COLORS)."""
        rgb = np.ascontiguousarray(rgb, np.float32).reshape(-1, 3)
        out = np.zeros(rgb.shape[0], np.uint32)
        self._ck(lib().pt_resolve_colors_noalloc(self._h, rgb.shape[0], _p(rgb), _p(out)))
        return out

    def measure_fp32_peak(self):
        v = C.c_float()
        self._ck(lib().pt_measure_fp32_peak(self._h, C.byref(v)))
        return v.value


def tile_ids(width, height, tile_rank, tile_count):
    """Row-major ids of the 32x32 tiles (src/win32_main.cpp:888-904) owned by tile_rank: tile (tx, ty) belongs to rank
    (tx + k * ty) % tile_count (pt_tile_owner, include/pt_api.h)."""
    from . import dist
    return dist.my_tiles(width, height, tile_rank, tile_count)
