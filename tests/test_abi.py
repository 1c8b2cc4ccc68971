"""The C-ABI libraries load and export every symbol their headers declare (no compute, no GPU)."""
import ctypes as C
import os
import re

from conftest import ROOT
from pathtracer_b200 import _ctypes_api as T


def _declared(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pt_[a-z0-9_]+)\s*\(", text)))


def test_pt_api_exports_every_declared_symbol():
    from pathtracer_b200 import render
    L = render.lib()
    names = _declared("pt_api.h")
    assert "pt_upload_scene" in names and "pt_render" in names and "pt_readback" in names
    for n in names:
        assert hasattr(L, n), n
    assert sorted(render.EXPORTS) == names
    assert L.pt_api_version() == 1


def test_release_library_reads_no_environment():
    """Experiment switches live in the debug build (make -C pathtracer_b200/csrc debug, -DPT_DEBUG_ENV); the library a
    host links holds no getenv of a PT_ name."""
    import subprocess
    from pathtracer_b200 import render
    out = subprocess.run(["strings", render.LIB_PATH], capture_output=True, text=True).stdout.splitlines()
    assert [s for s in out if re.fullmatch(r"PT_[A-Z0-9_]+", s)] == []


def test_pt_host_exports_every_declared_symbol():
    from pathtracer_b200 import host
    L = host.lib()
    for n in _declared("pt_host.h"):
        assert hasattr(L, n), n


def test_struct_sizes_match_header_contract():
    assert C.sizeof(T.Material) == 68 and C.sizeof(T.Plane) == 20 and C.sizeof(T.Quad) == 40
    assert C.sizeof(T.Sphere) == 20 and C.sizeof(T.Aabb) == 52 and C.sizeof(T.Camera) == 116
    assert C.sizeof(T.Scene) == 8 * 4 + 7 * 8 + 4 * 16 + 116
    assert T.HIT_DTYPE.itemsize == 28
    assert C.sizeof(T.Stats) == 56 and C.sizeof(T.Tuning) == 32 and C.sizeof(T.Config) == 24


def test_render_core_fails_loudly_without_a_device():
    """No CPU fallback: on a box without CUDA pt_create must return PT_ERR_NO_DEVICE."""
    import pytest
    from pathtracer_b200 import render
    try:
        rc = render.RenderCore(0)
    except render.PtError as e:
        assert e.code == -2 and "no CPU fallback" in str(e)
        return
    rc.close()
    pytest.skip("a CUDA device is present")


def test_product_does_not_import_the_oracle():
    """Nothing under pathtracer_b200/ (the product) may reference oracle/."""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "pathtracer_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                for pat in (r"^\s*(from|import)\s+oracle", r"#\s*include\s*[<\"].*oracle", r"libpt_oracle", r"libref_harness", r"pathtracer_ref"):
                    assert not re.search(pat, text, flags=re.M), (os.path.join(dirpath, f), pat)
