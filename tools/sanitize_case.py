"""Smallest program that runs every kernel of the library once or twice (for compute-sanitizer --tool memcheck):
batch extraction (staged-load + TMA K1, border, cell NMS, quadtree, describe<8>), single-frame extraction (latency variants),
stereo match (single + batch), batch SearchForInitialization (team kernel), windowed matchers, BoW, 2-NN (+ no-alloc), vocabulary."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from orb_slam_b200 import ORBextractor
from orb_slam_b200._lib import check
from orb_slam_b200.matcher import MatcherContext, ORBmatcher
from orb_slam_b200.synth import synth_frame, synth_descriptors
from tests import scenes
from tests.oracle_lib import OracleExtractor

w, h, nf, n = 752, 480, 1000, 24
frames = np.stack([synth_frame(w, h, 900, shift=(2 * i, -i)) for i in range(n)])
exL, exR = ORBextractor(nf, 1.2, 8, 20, 7), ORBextractor(nf, 1.2, 8, 20, 7)
ctx = MatcherContext(0)
cap = exL.max_keypoints(w, h)
bufs = []
for ex, shift in ((exL, 0), (exR, 9)):
    d_img = torch.from_numpy(np.roll(frames, -shift, axis=2)).cuda()
    d_k = torch.zeros((n, cap, 28), dtype=torch.uint8, device="cuda"); d_d = torch.zeros((n, cap, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(n, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    check(ex._L.orbx_extract_batch_device(ex.handle, d_img.data_ptr(), n, w, h, w, w * h, d_k.data_ptr(), d_d.data_ptr(), cap, d_n.data_ptr()), "batch")
    check(ex._L.orbx_synchronize(ex.handle), "sync")
    bufs.append((d_img, d_k, d_d, d_n))
d_u = torch.zeros((n, cap), dtype=torch.float32, device="cuda"); d_z = torch.zeros((n, cap), dtype=torch.float32, device="cuda")
d_nm = torch.zeros(n, dtype=torch.int32, device="cuda")
exL.stereo_match_batch_device(exR, n, bufs[0][1].data_ptr(), bufs[0][2].data_ptr(), bufs[0][3].data_ptr(), bufs[1][1].data_ptr(),
                              bufs[1][2].data_ptr(), bufs[1][3].data_ptr(), cap, 386.1448, 0.5371, d_u.data_ptr(), d_z.data_ptr(), d_nm.data_ptr())
exL.synchronize()
d_m = torch.zeros((n - 1, cap), dtype=torch.int32, device="cuda"); d_c = torch.zeros(n - 1, dtype=torch.int32, device="cuda")
ctx.search_for_initialization_batch_device(bufs[0][1].data_ptr(), bufs[0][2].data_ptr(), bufs[0][3].data_ptr(), cap, n, w, h,
                                           d_matches12=d_m.data_ptr(), d_nmatches=d_c.data_ptr())
ctx.synchronize()
print("batch: keypoints", int(bufs[0][3].sum()), "stereo", int(d_nm.sum()), "init matches", int(d_c.sum()))
# host-buffer pipeline (chunks) and the single-frame call
kl, dl = exL(frames[0]); kr, dr = exR(np.roll(frames[0], -9, axis=1))
u, z, nm = exL.stereo_match(exR, kl, dl, kr, dr, 386.1448, 0.5371)
F1, F2 = scenes.make_frame(kl, dl, w, h), scenes.make_frame(*exL(frames[1]), w, h)
prev = np.ascontiguousarray(np.stack([F1.mvKeysUn["x"], F1.mvKeysUn["y"]], 1), np.float32)
print("single: stereo", nm, "init", ORBmatcher(0.9, True, ctx).SearchForInitialization(F1, F2, prev, 100)[0])
pts = scenes.last_frame_points(F1, seed=2)
obs = np.full(F2.N, -1, np.int32)
print("projection", ORBmatcher(0.9, True, ctx).SearchByProjectionFrame(F2, F1, pts, obs, 15.0, True)[0])
s = scenes.bow_scene(seed=1)
print("bow", ORBmatcher(0.7, True, ctx).SearchByBoWFrame(s["desc1"], s["angle1"], s["valid1"], s["fv1"], s["desc2"], s["angle2"], s["fv2"])[0])
db = synth_descriptors(30001, 5); q = db[::97].copy()
d_db = torch.from_numpy(db).cuda()
r = ctx.bf2nn(q, d_db.data_ptr(), len(db))
d_q = torch.from_numpy(q).cuda(); d_o = torch.zeros(len(q) * 8, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
check(ctx._L.orbm_bf2nn_device_noalloc(ctx.h, d_q.data_ptr(), len(q), d_db.data_ptr(), len(db), 0, d_o.data_ptr()), "noalloc")
ctx.synchronize()
print("2nn ok", bool((r["i1"] == np.arange(0, 30001, 97)).all()), bool(d_o.cpu().numpy().tobytes() == r.tobytes()))
exL.close(); exR.close(); ctx.close()
print("SANITIZE_CASE_DONE")
