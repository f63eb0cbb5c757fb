"""Drives the two vocabulary-side kernels for an ncu capture: vocab_descend_kernel on a device-resident batch of 1.02 M
descriptors against a k = 10, L = 6 synthetic vocabulary, distinctive_kernel on 20 000 map points."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from orb_slam_b200._lib import check, lib
from orb_slam_b200.matcher import MatcherContext
from orb_slam_b200.synth import synth_vocabulary
from orb_slam_b200.vocabulary import ORBVocabulary, compute_distinctive_descriptors
from tests import scenes

ctx = MatcherContext(0)
arrays = synth_vocabulary(10, 6)
voc = ORBVocabulary.from_arrays(ctx, 10, 6, *arrays)
q = scenes.vocabulary_queries(arrays[2], arrays[1], 8000, seed=1)
big = torch.from_numpy(q).cuda().repeat(128, 1).contiguous()          # 1 024 000 descriptors
res = torch.zeros((big.shape[0], 2), dtype=torch.int32, device="cuda")
torch.cuda.synchronize()
L = lib()
for rep in range(5):
    t0 = time.perf_counter()
    check(L.orbm_vocab_descend_device(ctx.h, voc._h, big.data_ptr(), big.shape[0], 4, res.data_ptr()), "descend")
    check(L.orbm_synchronize(ctx.h), "sync")
    print("descend %d features: %.3f ms" % (big.shape[0], (time.perf_counter() - t0) * 1e3))
D, start = scenes.observation_lists(20000, seed=5, max_obs=10, big=(40, 77, 150))
for rep in range(3):
    t0 = time.perf_counter()
    compute_distinctive_descriptors(ctx, D, start)
    print("distinctive %d points / %d observations: %.3f ms" % (len(start) - 1, len(D), (time.perf_counter() - t0) * 1e3))
