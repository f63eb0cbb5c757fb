"""Drives the batched SearchForInitialization (BASELINE.json configs[2]) for timing / ncu: 1024 EuRoC-size frames extracted on
the device, 1023 consecutive pairs matched; prints CUDA-event times of the two matcher kernels."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from orb_slam_b200 import ORBextractor
from orb_slam_b200._lib import check
from orb_slam_b200.matcher import MatcherContext
from orb_slam_b200.synth import synth_frame

w, h, nf, distinct, nframes = 752, 480, 1000, 32, 1024
base = np.stack([synth_frame(w, h, 900, shift=(2 * i, -i)) for i in range(distinct)])
order = [(i % (2 * distinct - 2)) for i in range(nframes)]
order = [j if j < distinct else 2 * distinct - 2 - j for j in order]
d_img = torch.from_numpy(base).cuda()[torch.tensor(order, device="cuda")].contiguous()
ex = ORBextractor(nf, 1.2, 8, 20, 7, device=0)
ctx = MatcherContext(0)
cap = ex.max_keypoints(w, h)
d_k = torch.zeros((nframes, cap, 28), dtype=torch.uint8, device="cuda"); d_d = torch.zeros((nframes, cap, 32), dtype=torch.uint8, device="cuda")
d_n = torch.zeros(nframes, dtype=torch.int32, device="cuda")
d_m = torch.zeros((nframes - 1, cap), dtype=torch.int32, device="cuda"); d_nm = torch.zeros(nframes - 1, dtype=torch.int32, device="cuda")
torch.cuda.synchronize()
check(ex._L.orbx_extract_batch_device(ex.handle, d_img.data_ptr(), nframes, w, h, w, w * h, d_k.data_ptr(), d_d.data_ptr(), cap, d_n.data_ptr()), "extract")
check(ex._L.orbx_synchronize(ex.handle), "sync")
stream = torch.cuda.ExternalStream(ctx._L.orbm_stream(ctx.h))
for rep in range(4):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    ctx.search_for_initialization_batch_device(d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(), cap, nframes, w, h,
                                               d_matches12=d_m.data_ptr(), d_nmatches=d_nm.data_ptr())
    e1.record(stream)
    ctx.synchronize()
    print("grid + match: %.3f ms, mean matches %.1f" % (e0.elapsed_time(e1), d_nm.float().mean().item()))
