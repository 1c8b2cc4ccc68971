"""The C oracle against the reference ITSELF (oracle/_ref, built from /root/reference by `make -C oracle ref`),
live, on inputs the committed fixtures do not cover: other seeds, sizes, the thin lens, random rays, the texture
flags, the whole RenderTexel path down to the BGRA bytes.  Skipped where oracle/_ref was never built.
Everything is BIT-EXACT: same mt19937 draws in the same order, same float operations in the same order."""
import numpy as np
import pytest

from oracle import pt_oracle as O
from pathtracer_b200 import _ctypes_api as T


@pytest.mark.parametrize("w,W,H", [(1, 96, 54), (2, 96, 54), (3, 80, 60), (4, 96, 54), (5, 128, 72)])
def test_random_rays_bit_exact(ref, worlds, w, W, H):
    ref.load(w, W, H, 2, 0, 7)
    sc = worlds(w).scene(W, H)
    rng = np.random.default_rng(40 + w)
    n = 20000
    cam = np.array([sc.camera.pos.x, sc.camera.pos.y, sc.camera.pos.z], np.float32)
    tgt = np.array([sc.camera.target.x, sc.camera.target.y, sc.camera.target.z], np.float32)
    scale = np.float32(np.linalg.norm(cam - tgt))
    o = (tgt + (rng.random((n, 3)) - 0.5) * scale).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    a, b = ref.intersect(o, d), O.intersect(sc, o, d)
    assert np.array_equal(a["t"], b["t"]) and np.array_equal(a["mat"], b["mat"])
    assert np.array_equal(a["kind"], b["kind"]) and np.array_equal(a["index"], b["index"])
    assert np.array_equal(a["normal"].view(np.uint32), b["normal"].view(np.uint32))
    ref.seed(77)
    ra = ref.raycast(o[:4000], d[:4000])
    O.seed_mt(77)
    rb = O.raycast(sc, o[:4000], d[:4000], O.RNG_MT19937)
    assert np.array_equal(ra.view(np.uint32), rb.view(np.uint32))


@pytest.mark.parametrize("w,lens,flags,pp,seed", [(1, False, 0, 3, 5), (1, False, T.SCENE_NORMAL_MAPS | T.SCENE_METALNESS_MAPS, 2, 6),
                                                  (2, True, 7, 2, 7), (3, True, 7, 3, 8), (4, False, 7, 2, 9), (5, True, 7, 3, 10),
                                                  (5, False, 7, 5, 11)])
def test_render_texel_bit_exact(ref, worlds, w, lens, flags, pp, seed):
    """RenderTexel's two sampling loops + RayCast + tonemap + pack: HDR sums and LDR bytes of a window."""
    W, H = 64, 40
    # harness flag bits: 1 = -d (thin lens), 2 = -n, 4 = -m, 8 = -r (disable normal / metalness / roughness maps)
    hflags = (1 if lens else 0) | (0 if flags & T.SCENE_NORMAL_MAPS else 2) | (0 if flags & T.SCENE_METALNESS_MAPS else 4)
    hflags |= 0 if flags & T.SCENE_ROUGHNESS_MAPS else 8
    ref.load(w, W, H, pp, hflags, 7)
    sc = worlds(w).scene(W, H, lens=lens, tex_flags=flags)
    ref.seed(seed)
    a = ref.render_hdr(8, 4, 40, 30)
    O.seed_mt(seed)
    b = O.render(sc, W, H, pp, O.RNG_MT19937, window=(8, 4, 40, 30))
    assert np.array_equal(a["sum"], b["sum"]) and np.array_equal(a["sumsq"], b["sumsq"])
    assert np.array_equal(a["ldr"], b["ldr"])


def test_resolve_bit_exact(ref):
    rng = np.random.default_rng(1)
    cols = np.concatenate([rng.random((3000, 3)) ** 3 * 6, rng.random((1000, 3)) * 0.004]).astype(np.float32)
    assert np.array_equal(ref.resolve(cols), O.resolve(cols))
