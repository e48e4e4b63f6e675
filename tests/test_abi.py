"""CPU tests: the C-ABI library loads and exports every symbol include/mi_b200.h declares
(no compute calls without a GPU), and it refuses to run without a CUDA device."""
import ctypes as C
import os
import re

import pytest

from mitsuba3_nasa_b200 import _abi as A, _lib, scene as S, scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _built():
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__ as g
        g.build()


def test_library_exports_every_declared_symbol():
    _built()
    l = _lib.lib()
    header = open(os.path.join(ROOT, "include", "mi_b200.h")).read()
    declared = set(re.findall(r"\b(mib_[a-z0-9_]+)\s*\(", header))
    assert declared == set(A.EXPORTS), declared ^ set(A.EXPORTS)
    for name in declared:
        assert hasattr(l, name), name


def test_struct_layout_matches_header():
    _built()
    l = _lib.lib()
    l.mib_sizeof_scene_desc.restype = C.c_size_t
    assert l.mib_sizeof_scene_desc() == C.sizeof(A.SceneDesc)
    assert l.mib_abi_version() == A.MIB_ABI_VERSION


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    _built()
    sc = S.load_dict(scenes.c1_homogeneous_cube(res=4, spp=1), "llvm_ad_rgb")
    with pytest.raises(RuntimeError, match="no CUDA device|CUDA"):
        _lib.DeviceScene(sc, 0)
