"""CPU: the product library loads and exports every function include/*.h declares (no compute calls)."""
import ctypes
import glob
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    names = []
    for hdr in sorted(glob.glob(os.path.join(ROOT, "include", "orb*.h"))):
        if hdr.endswith("orb_pattern_31.h"):
            continue
        src = re.sub(r"/\*.*?\*/", "", open(hdr).read(), flags=re.S)
        names += re.findall(r"\b(orb[xm]_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))


def test_library_exports_every_declared_symbol():
    from orb_slam_b200 import LIB_PATH
    assert os.path.exists(LIB_PATH), "build first: python -c 'import __graft_entry__ as g; g.build()'"
    lib = ctypes.CDLL(LIB_PATH)
    names = declared_functions()
    assert len(names) >= 15
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_binding_table_loads():
    from orb_slam_b200._lib import lib
    lib()


def test_keypoint_layout():
    from orb_slam_b200 import KEYPOINT_DTYPE
    assert KEYPOINT_DTYPE.itemsize == 28
    assert [KEYPOINT_DTYPE.fields[n][1] for n in ("x", "y", "size", "angle", "response", "octave", "class_id")] == \
        [0, 4, 8, 12, 16, 20, 24]


def test_no_product_import_of_oracle():
    """the product package must never load, link or include anything under oracle/"""
    bad = re.compile(r"liborb_oracle|liborb_ref|oracle_lib|#include\s*[\"<][^\">]*oracle|from\s+tests|import\s+tests|orbo_|orbref_|cvprim")
    for path in glob.glob(os.path.join(ROOT, "orb_slam_b200", "**", "*"), recursive=True):
        if os.path.isfile(path) and path.endswith((".py", ".cu", ".cuh", ".h", ".cc", ".cpp", "Makefile")):
            txt = open(path, errors="ignore").read()
            assert not bad.search(txt), path
