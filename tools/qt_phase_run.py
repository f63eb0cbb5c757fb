"""Phase clocks of the level-0 quadtree block (debug build with -DQT_PROFILE, see tools/_prof): single frame and batch."""
import ctypes as C, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from orb_slam_b200 import ORBextractor
from orb_slam_b200._lib import check
from orb_slam_b200.synth import synth_frame
W, H = 1241, 376
ex = ORBextractor(2000, 1.2, 8, 20, 7); L = ex._L; h = ex.handle; cap = ex.max_keypoints(W, H)
L.orbx_extract_batch_device.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
L.orbx_synchronize.argtypes = [C.c_void_p]
for B in (1, 256):
    dimg = torch.zeros((B, H, 1280), dtype=torch.uint8, device="cuda")
    for i in range(min(B, 8)):
        dimg[i::8, :, :W] = torch.from_numpy(synth_frame(W, H, i)).cuda()
    dk = torch.empty((B, cap, 28), dtype=torch.uint8, device="cuda"); dd = torch.empty((B, cap, 32), dtype=torch.uint8, device="cuda")
    dn = torch.zeros(B + 4, dtype=torch.int32, device="cuda")
    for r in range(3):
        check(L.orbx_extract_batch_device(h, dimg.data_ptr(), B, W, H, 1280, 1280 * H, dk.data_ptr(), dd.data_ptr(), cap, dn.data_ptr()), "dev")
        check(L.orbx_synchronize(h), "sync")
    c = np.zeros(64, np.int64)
    L.orbx_debug_qt_clock(C.c_void_p(c.ctypes.data))
    us = lambda a, b: (c[b] - c[a]) / 1965.0
    print(f"B={B}: M={c[39]} iterations={c[7]}  gather {us(0,1):.1f}  codes {us(1,2):.1f}  sort {us(2,3):.1f}  init {us(3,4):.1f}  loop {us(4,5):.1f}  select {us(5,6):.1f}  total {us(0,6):.1f} us")
    it = int(c[7])
    print("   passes (n at entry, negative = priority pass : us):", [(int(c[40 + k]), round(float((c[8 + k + 1] if k + 1 < it else c[5]) - c[8 + k]) / 1965.0, 1)) for k in range(min(it, 20))])
