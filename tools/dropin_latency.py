"""Wall time of the C++ drop-in ORBmatcher members (orb_slam_b200/csrc/host) as tests/cpp/dropin_harness.cc calls them:
the parity tests of tests/test_dropin_cpp_gpu.py are run a few times each and the harness's own timer around the class
call (conversion of Frame / KeyFrame / MapPoint state + C ABI + device work) is read after every call."""
import ctypes as C, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import test_dropin_cpp_gpu as T


class Rec:
    def __init__(self, lib):
        self.lib, self.ms = lib, {}
        lib.dropin_last_ms.restype = C.c_double

    def __getattr__(self, name):
        f = getattr(self.lib, name)
        if not name.startswith("dropin_"):
            return f

        def call(*a):
            r = f(*a)
            self.ms.setdefault(name, []).append(self.lib.dropin_last_ms())
            return r
        return call


rec = Rec(C.CDLL(T.SO))
CASES = ["test_search_by_projection_members", "test_search_by_bow_members", "test_search_by_projection_keyframe_member",
         "test_search_for_triangulation_member", "test_fuse_member", "test_search_by_projection_sim3_member", "test_fuse_sim3_member",
         "test_search_by_sim3_member"]
for name in CASES:
    fn = getattr(T, name)
    params = ()
    for mk in getattr(fn, "pytestmark", []):
        if mk.name == "parametrize":
            first = mk.args[1][0]
            params = tuple(first) if isinstance(first, (tuple, list)) else (first,)
    for _ in range(4):
        fn(rec, *params)
for k, v in rec.ms.items():
    print(f"{k:44s} first {v[0]:7.3f} ms   steady (min of {len(v) - 1}) {min(v[1:]):7.3f} ms")
