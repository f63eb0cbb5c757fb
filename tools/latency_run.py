"""Single-frame latency of orbx_extract (pageable and pinned image), with the ORBX_TRACE device timeline."""
import ctypes as C, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from orb_slam_b200 import ORBextractor
from orb_slam_b200._lib import check
from orb_slam_b200.synth import synth_frame
W, H = 1241, 376
ex = ORBextractor(2000, 1.2, 8, 20, 7); L = ex._L; h = ex.handle; cap = ex.max_keypoints(W, H)
img = synth_frame(W, H, 3)
pin = torch.empty((H, W), dtype=torch.uint8, pin_memory=True); pin.numpy()[:] = img
k = np.empty((cap, 28), np.uint8); d = np.empty((cap, 32), np.uint8); n = C.c_int()
kp = torch.empty((cap, 28), dtype=torch.uint8, pin_memory=True); dp = torch.empty((cap, 32), dtype=torch.uint8, pin_memory=True)
for name, ptr, ko, do in (("pageable", img.ctypes.data, k.ctypes.data, d.ctypes.data), ("pinned", pin.data_ptr(), kp.data_ptr(), dp.data_ptr())):
    lat = []
    for i in range(200):
        t0 = time.perf_counter()
        check(L.orbx_extract(h, ptr, W, H, W, ko, do, cap, C.byref(n)), "orbx_extract")
        lat.append(time.perf_counter() - t0)
    lat = np.array(lat[20:]) * 1e3
    print(f"{name}: median {np.median(lat):.3f} ms  p10 {np.percentile(lat, 10):.3f}  p90 {np.percentile(lat, 90):.3f}  n={n.value}")
# per-stage device time of a single frame (profiling mode: stages run back to back with events in between)
L.orbx_set_profiling.argtypes = [C.c_void_p, C.c_int]; L.orbx_stage_times.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
L.orbx_extract_batch_device.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
L.orbx_synchronize.argtypes = [C.c_void_p]
dimg = torch.zeros((H, 1280), dtype=torch.uint8, device="cuda"); dimg[:, :W] = torch.from_numpy(img).cuda()
dk = torch.empty((cap, 28), dtype=torch.uint8, device="cuda"); dd = torch.empty((cap, 32), dtype=torch.uint8, device="cuda")
dn = torch.zeros(4, dtype=torch.int32, device="cuda")
torch.cuda.synchronize()
check(L.orbx_set_profiling(h, 1), "prof")
acc = np.zeros(5)
t = np.zeros(5, np.float32)
for i in range(50):
    check(L.orbx_extract_batch_device(h, dimg.data_ptr(), 1, W, H, 1280, 1280 * H, dk.data_ptr(), dd.data_ptr(), cap, dn.data_ptr()), "dev")
    check(L.orbx_synchronize(h), "sync")
    check(L.orbx_stage_times(h, t.ctypes.data, 5), "st")
    if i >= 10:
        acc += t
print("single-frame stage ms:", dict(zip(["pyramid_blur_fast", "border", "cell_nms", "quadtree", "describe"], (acc / 40).round(4).tolist())))
# the stereo constructor's pattern through the C++ drop-in class: two ORBextractor instances on two std::threads
# (reference src/Frame.cc:94-103), mvImagePyramid downloaded as the class does by default; per-pair time from the
# difference of a long and a short run (constructors excluded)
harness = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "cpp", "libdropin_harness.so")
if os.path.exists(harness):
    Hn = C.CDLL(harness)
    right = synth_frame(W, H, 3, shift=(-20, 0))
    kl = np.empty((cap, 28), np.uint8); dl = np.empty((cap, 32), np.uint8); kr = np.empty((cap, 28), np.uint8); dr = np.empty((cap, 32), np.uint8)
    nl, nr = C.c_int(), C.c_int()
    def run(reps):
        t0 = time.perf_counter()
        Hn.dropin_extract_stereo(C.c_void_p(img.ctypes.data), C.c_void_p(right.ctypes.data), W, H, 2000, C.c_void_p(kl.ctypes.data), C.c_void_p(dl.ctypes.data),
                                 C.byref(nl), C.c_void_p(kr.ctypes.data), C.c_void_p(dr.ctypes.data), C.byref(nr), cap, reps)
        return time.perf_counter() - t0
    run(3)
    a = min(run(5) for _ in range(3)); b = min(run(205) for _ in range(3))
    print(f"C++ drop-in stereo pair (2 std::threads spawned per pair, pyramids downloaded): {(b - a) / 200 * 1e3:.3f} ms per pair, n = {nl.value}/{nr.value}")
    me, mm, nm = C.c_double(), C.c_double(), C.c_int()
    for dl_ in (1, 0):
        Hn.dropin_time_stereo_frame(C.c_void_p(img.ctypes.data), C.c_void_p(right.ctypes.data), W, H, 2000, C.c_float(386.1448), C.c_float(0.537),
                                    dl_, 200, C.byref(me), C.byref(mm), C.byref(nm))
        print(f"C++ stereo Frame constructor, pyramids {'downloaded' if dl_ else 'kept on the device'}: ExtractORB x2 {me.value:.3f} ms + ComputeStereoMatches {mm.value:.3f} ms per pair ({nm.value} stereo matches)")
# do two single-frame chains overlap on the device?  (two extractors, device-resident input, enqueued from one thread)
exB = ORBextractor(2000, 1.2, 8, 20, 7); hB = exB.handle
dk2 = torch.empty((cap, 28), dtype=torch.uint8, device="cuda"); dd2 = torch.empty((cap, 32), dtype=torch.uint8, device="cuda"); dn2 = torch.zeros(4, dtype=torch.int32, device="cuda")
check(L.orbx_set_profiling(h, 0), "prof")
def dev(hh, k_, d_, n_):
    check(L.orbx_extract_batch_device(hh, dimg.data_ptr(), 1, W, H, 1280, 1280 * H, k_.data_ptr(), d_.data_ptr(), cap, n_.data_ptr()), "dev")
for both in (False, True):
    ts = []
    for i in range(60):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        dev(h, dk, dd, dn)
        if both:
            dev(hB, dk2, dd2, dn2)
        L.orbx_synchronize(h)
        if both:
            L.orbx_synchronize(hB)
        ts.append(time.perf_counter() - t0)
    print(f"device-resident single-frame chain{'s x2 on two instances' if both else ''}: {np.median(ts[10:]) * 1e3:.3f} ms")
