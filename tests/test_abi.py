"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/sirrt_b200.h declares, and fails loudly (no fallback) when no CUDA device is present."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "sirrt_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sirrt_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(pkg):
    lib = ctypes.CDLL(pkg.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/sirrt_b200.h but not exported"
    assert set(names) == set(pkg.EXPORTS), set(names) ^ set(pkg.EXPORTS)


def test_no_cpu_fallback_without_a_device(pkg, packs):
    if pkg.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(pkg.SirrtError) as e:
        pkg.Scene(packs("synA_8_50"))
    assert e.value.code == -3 and "no CPU fallback" in str(e.value)
    with pytest.raises(pkg.SirrtError):
        pkg.KnnIndex(6, 100)


def test_product_never_references_the_oracle():
    """The product path must not import, link or call anything under oracle/."""
    pkg_dir = os.path.join(ROOT, "manipulationplanning-si-rrt_b200")
    for d, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".inl", ".cpp", ".h")):
                txt = open(os.path.join(d, f), errors="replace").read()
                assert "sirrt_oracle" not in txt and "oracle_" not in txt and "load_oracle" not in txt, os.path.join(d, f)


def test_bad_pack_is_rejected_before_any_cuda_call(pkg):
    h = ctypes.c_void_p()
    lib = ctypes.CDLL(pkg.LIB_PATH)
    rc = lib.sirrt_scene_create_from_pack(None, 0, 0, 0, ctypes.byref(h))
    assert rc == -1


def test_option_constants_match_the_header():
    """The Python mirror's OPT_* values are the SIRRT_OPT_* enumerators of include/sirrt_b200.h (the tests toggle kernels paths
    through them: a drifted value would silently toggle another knob)."""
    import re

    g = __import__("__graft_entry__")
    hdr = open(os.path.join(g.ROOT, "include", "sirrt_b200.h")).read()
    enum = dict((m.group(1), int(m.group(2))) for m in re.finditer(r"SIRRT_OPT_([A-Z_0-9]+)\s*=\s*(\d+)", hdr))
    src = open(os.path.join(g.PKG_DIR, "__init__.py")).read()
    m = re.search(r"^(OPT_[A-Z_, ]+?)\s*=\s*([\d, ]+)$", src, re.M)
    names = [n.strip()[4:] for n in m.group(1).split(",")]
    values = [int(v) for v in m.group(2).split(",")]
    assert len(names) == len(values) == len(enum), (names, enum)
    for n, v in zip(names, values):
        assert enum[n] == v, (n, v, enum[n])
