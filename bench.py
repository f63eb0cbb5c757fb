#!/usr/bin/env python
"""bench.py — ORB frames/s on the KITTI-stereo workload (BASELINE.json configs[1]).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU path (oracle/_ref)

A step = one pass of ORBextractor::operator() over one batch of synthetic
1241x376 frames (left/right stereo pairs, nFeatures 2000, 8 levels, 1.2, FAST
20/7).  `value` = frames/s with the batch resident in HBM; `e2e` = frames/s
through the host-buffer C ABI call (pinned host images in, keypoints +
descriptors out, both copies inside the timed region).  One JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W_, H_, NFEAT = 1241, 376, 2000
SETTINGS = dict(nFeatures=NFEAT, scaleFactor=1.2, nLevels=8, iniThFAST=20, minThFAST=7)
METRIC = "ORB frames/s (1241x376, 2000 kp)"
UNIT = "frames/s"
STAGES = ["pyramid_blur_fast", "border", "cell_nms", "quadtree", "describe"]


def level_sizes(w, h, nlevels=8, sf=1.2):
    s = np.float32(1.0)
    out = []
    for _ in range(nlevels):
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float32(w) * inv)), int(np.rint(np.float32(h) * inv))))
        s = np.float32(np.float64(s) * np.float64(np.float32(sf)))
    return out


def algorithmic_bytes(w, h, nkp):
    """B_alg = W*H + A' + A + N*60, split per stage."""
    lv = level_sizes(w, h)
    A = sum(a * b for a, b in lv)
    Ab = sum((a + 38) * (b + 38) for a, b in lv)
    per_stage = {"pyramid_blur_fast": w * h + 2 * A, "border": Ab - A, "describe": nkp * 60}
    return w * h + Ab + A + nkp * 60, per_stage


def make_frames(n_distinct, start):
    """left/right pairs: right = left scene shifted by a per-pair disparity"""
    from orb_slam_b200.synth import synth_frame
    frames = []
    for i in range(n_distinct // 2):
        idx = start + i
        frames.append(synth_frame(W_, H_, idx))
        frames.append(synth_frame(W_, H_, idx, shift=(-(5 + (idx * 7) % 56), 0)))
    return np.stack(frames)


def ncu_pipe_utilisation():
    """ALU-pipe and issue-slot utilisation of the dominant kernel from the committed ncu capture (profiles/traffic.json): the
    figures that say how close the ALU-bound K1 is to its real bound."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        k = t["kernels"]["pyr_level_kernel"]
        return {"alu_pipe_pct": round(k["alu_pipe_pct"], 1), "issue_active_pct": round(k["issue_active_pct"], 1), "source": "profiles/" + t["source"].replace(".ncu-rep", "") + " (ncu --set full)"}
    except Exception:
        return None


def measured_traffic(batch):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel's launches (all 8 levels of pyr_level_kernel)
    from the committed `ncu --set full` capture, scaled from the capture's batch to this run's (profiles/*.json)."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        return t["pyramid_blur_fast_dram_bytes_per_frame"] * batch
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.lines = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 6:
                continue
            try:
                sm.append(float(p[0])); mx = float(p[1])
            except ValueError:
                continue
            for nm, v in zip(names, p[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_extract_throughput(frames, n_threads, frames_per_thread, kind):
    """frames/s of the CPU path with one extractor instance per thread (the reference's own threading
    model, src/Frame.cc:94-103).  kind 'reference' = oracle/_ref (the reference's ORBextractor.cc,
    system allocator), 'port' = oracle/liborb_oracle.so."""
    from tests import oracle_lib
    if kind == "reference":
        mk = lambda: oracle_lib.RefExtractor(NFEAT, 1.2, 8, 20, 7)
        call = lambda ex, im: ex(im, canonical=False)
    else:
        mk = lambda: oracle_lib.OracleExtractor(NFEAT, 1.2, 8, 20, 7)
        call = lambda ex, im: ex(im)
    exs = [mk() for _ in range(n_threads)]
    for ex in exs:
        call(ex, frames[0])          # warm-up (page faults, tables)
    barrier = threading.Barrier(n_threads + 1)

    def work(t):
        barrier.wait()
        for k in range(frames_per_thread):
            call(exs[t], frames[(t * frames_per_thread + k) % len(frames)])
        barrier.wait()

    th = [threading.Thread(target=work, args=(t,)) for t in range(n_threads)]
    for t in th:
        t.start()
    barrier.wait()
    t0 = time.perf_counter()
    barrier.wait()
    dt = time.perf_counter() - t0
    for t in th:
        t.join()
    return n_threads * frames_per_thread / dt, dt


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU ORBextractor on the host cores (rank 0 only)."""
    if rank != 0:
        return
    from tests import oracle_lib
    kind = "reference" if oracle_lib.have_ref() else "port"
    cores = os.cpu_count() or 1
    frames = make_frames(16, 0)
    per_thread = max(2, args.cpu_frames // cores)
    vals = []
    for _ in range(args.warmup):
        cpu_extract_throughput(frames, cores, 1, kind)
    t_all = 0.0
    for _ in range(args.steps):
        v, dt = cpu_extract_throughput(frames, cores, per_thread, kind)
        vals.append(v); t_all += dt
    value = cores * per_thread * args.steps / t_all
    single, _ = cpu_extract_throughput(frames, 1, 8, kind)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t_all / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "KITTI-stereo 1241x376 left+right, nFeatures 2000, 8 levels, 1.2, FAST 20/7", **SETTINGS,
                   "frames_per_step": cores * per_thread},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": f"{cores} threads x {per_thread} frames per step, one extractor instance per thread",
                         "single_thread_frames_per_s": single},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


HAM_NQ = 4096
HAM_ROWS_PER_GPU = 1_250_000          # BASELINE.json configs[3]: 10M rows over 8 GPUs


def make_hamming_case(rank, world):
    """config 4: uniform random database (seed 0xDB, counter-based so every rank can build its own shard),
    queries = database rows at stride 2441 with up to 40 flipped bits (identical on every rank)."""
    from orb_slam_b200.sharded import shard_range
    from orb_slam_b200.synth import synth_descriptors, splitmix64
    ndb = HAM_ROWS_PER_GPU * world
    lo, hi = shard_range(ndb, world, rank)
    # rows are generated from the global row counter: shard = slice of the global stream (32 B = 4 words per row)
    idx = np.arange(lo * 4, hi * 4, dtype=np.uint64)
    with np.errstate(over="ignore"):
        shard = splitmix64(splitmix64(np.uint64(0xDB)) ^ idx).view(np.uint8).reshape(-1, 32)
    qrows = (np.arange(HAM_NQ, dtype=np.uint64) * np.uint64(2441)) % np.uint64(ndb)
    qi = (qrows[:, None] * np.uint64(4) + np.arange(4, dtype=np.uint64)[None, :]).ravel()
    with np.errstate(over="ignore"):
        q = splitmix64(splitmix64(np.uint64(0xDB)) ^ qi).view(np.uint8).reshape(-1, 32).copy()
    rng = np.random.default_rng(7)
    for i in range(HAM_NQ):
        for b in rng.integers(0, 256, rng.integers(0, 41)):
            q[i, b >> 3] ^= 1 << (b & 7)
    return shard, lo, q, ndb


def run_hamming(args, rank, world, local_rank, dist):
    """Sharded brute-force Hamming 2-NN: per-rank scan + NCCL all-gather + exact merge."""
    import torch
    from orb_slam_b200.matcher import TOP2_DTYPE
    from orb_slam_b200.sharded import ShardedDatabase
    dev = torch.device("cuda", local_rank)
    shard, lo, q, ndb = make_hamming_case(rank, world)
    sd = ShardedDatabase(shard, lo, device=local_rank)
    d_q = torch.from_numpy(q).to(dev)
    steps, warm = max(3, args.steps), max(3, args.warmup)
    for _ in range(warm):
        sd.query_device(d_q)
    sd.stream.synchronize(); torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(sd.stream)
    for _ in range(steps):
        out = sd.query_device(d_q)
    e1.record(sd.stream)
    sd.stream.synchronize(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    # end to end: host queries in, host top-2 out
    t0 = time.perf_counter()
    for _ in range(steps):
        res = sd.query(q)
    e2e = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX); dist.all_reduce(e2e, op=dist.ReduceOp.MAX)
    if rank != 0:
        return None
    popc_peak = sd.ctx.popc_peak()
    sec = float(ms.item()) * 1e-3 / steps
    pairs = HAM_NQ * float(ndb)
    # analytic check: a query with k <= 40 flips of row r has distance <= k to row r, far below any random row (~128 +- 8)
    exp_rows = (np.arange(HAM_NQ, dtype=np.int64) * 2441) % ndb
    ok = bool((res["i1"].astype(np.int64) == exp_rows).all() and (res["d1"] <= 40).all() and (res["d2"] > 60).all())
    return {"metric": "Hamming 2-NN queries/s", "value": HAM_NQ / sec, "unit": "queries/s", "ms_per_step": sec * 1e3,
            "config": {"workload": "BASELINE.json configs[3] at %d GPU(s): %d queries vs %d-row database (%d rows per GPU), "
                                   "NCCL all-gather top-2 merge" % (world, HAM_NQ, ndb, HAM_ROWS_PER_GPU), "scaling": "weak"},
            "e2e": {"value": HAM_NQ * steps / float(e2e.item()), "unit": "queries/s", "h2d_bytes_per_step": HAM_NQ * 32,
                    "d2h_bytes_per_step": HAM_NQ * 8},
            "roofline": {"bound": "int-popc", "achieved": pairs * 8 / sec / 1e12 / world, "peak": popc_peak / 1e12,
                         "unit": "Tpopc32/s per GPU", "frac": pairs * 8 / sec / world / popc_peak,
                         "peak_source": "orbm_popc_peak microbenchmark on this GPU",
                         "note": "achieved counts the 8 algorithmic popc32 per pair (DescriptorDistance); the kernel executes 4 "
                                 "after carry-save compression, so frac can exceed 1",
                         "hbm_GBps_per_gpu": HAM_ROWS_PER_GPU * 32 * ((HAM_NQ + 1023) // 1024) / sec / 1e9},
            "pairs_per_s_all_gpus": pairs / sec,
            "nearest_neighbours_verified": ok}


def cpu_hamming_baseline(cores):
    """DescriptorDistance loop (oracle port of src/ORBmatcher.cc:1728-1744 + the top-2 scan) on a bounded sample"""
    from tests.oracle_lib import OracleMatcher
    from orb_slam_b200.synth import synth_descriptors
    db = synth_descriptors(HAM_ROWS_PER_GPU, 0xDB)
    per = 16
    q = db[: cores * per].copy()
    o = OracleMatcher()
    o.bf2nn(q[:1], db[:1000])
    barrier = threading.Barrier(cores + 1)

    def work(t):
        barrier.wait()
        o.bf2nn(q[t * per:(t + 1) * per], db)
        barrier.wait()
    th = [threading.Thread(target=work, args=(t,)) for t in range(cores)]
    for t in th:
        t.start()
    barrier.wait(); t0 = time.perf_counter(); barrier.wait(); dt = time.perf_counter() - t0
    for t in th:
        t.join()
    return {"value": cores * per / dt, "unit": "queries/s", "cores": cores, "kind": "port",
            "sample": f"{cores} threads x {per} queries vs {HAM_ROWS_PER_GPU} rows ({dt:.1f} s wall)"}


def run_stereo_frontend(local_rank, with_cpu):
    """BASELINE.json configs[1] as Tracking runs it, one stereo pair at a time (latency, not throughput): ExtractORB on
    two extractor instances from two host threads (src/Frame.cc:94-103), Frame::ComputeStereoMatches, then
    ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th=7) with 2000 synthetic map points.  Median ms per pair."""
    from orb_slam_b200 import ORBextractor
    from orb_slam_b200.matcher import MatcherContext, ORBmatcher
    from orb_slam_b200.synth import synth_frame
    from tests import scenes
    exL, exR = ORBextractor(NFEAT, 1.2, 8, 20, 7, device=local_rank), ORBextractor(NFEAT, 1.2, 8, 20, 7, device=local_rank)
    ctx = MatcherContext(local_rank)
    mbf = float(scenes.KITTI_BF); mb = float(np.float32(scenes.KITTI_BF) / scenes.KITTI_K[0, 0])
    # one scene drifting 2 px per frame (so consecutive frames do match), right image = left shifted by the disparity
    pairs = [(synth_frame(W_, H_, 500, shift=(-2 * i, 0)), synth_frame(W_, H_, 500, shift=(-2 * i - 15, 0))) for i in range(6)]
    out = {}

    from concurrent.futures import ThreadPoolExecutor
    pool = ThreadPoolExecutor(max_workers=1)             # the second host thread of src/Frame.cc:94-103, kept alive between frames

    def extract_pair(l, r, a, b):
        fut = pool.submit(b, r)
        left = a(l)
        return [left, fut.result()]

    def one(l, r, a, b, stereo, proj, last):
        t0 = time.perf_counter()
        (kl, dl), (kr, dr) = extract_pair(l, r, a, b)
        t1 = time.perf_counter()
        u, z, n = stereo(kl, dl, kr, dr)
        t2 = time.perf_counter()
        cur = scenes.make_frame(kl, dl, W_, H_, stereo=False, Tcw=np.eye(4, dtype=np.float32))
        cur.mvuRight = u
        nm = 0
        t3 = time.perf_counter()
        if last is not None:
            nm = proj(cur, *last)
        t4 = time.perf_counter()
        return (t1 - t0, t2 - t1, t4 - t3), (cur, kl, dl), n, nm

    def gpu_stereo(kl, dl, kr, dr):
        return exL.stereo_match(exR, kl, dl, kr, dr, mbf, mb)

    def featvec(d):
        # synthetic vocabulary: a feature's node is a hash of a few descriptor bits (no ORBvoc file on this box)
        from orb_slam_b200.matcher import FeatureVector
        nodes = {}
        for i, nd in enumerate(((d[:, 0].astype(np.uint32) ^ (d[:, 7].astype(np.uint32) << 1)) % np.uint32(100)).tolist()):
            nodes.setdefault(nd, []).append(i)
        return FeatureVector(nodes)

    def bow_inputs(cur, last):
        (kc, dc), (kl, dl) = cur, last
        valid = (np.arange(len(dl)) % 5 != 0).astype(np.uint8)                 # 80 % of the key frame's features hold a map point
        return dl, kl["angle"].astype(np.float32), valid, featvec(dl), dc, kc["angle"].astype(np.float32), featvec(dc)

    def gpu_proj(cur, lastF, pts, obs):
        return ORBmatcher(0.9, True, ctx).SearchByProjectionFrame(cur, lastF, pts, obs, 7.0, False)[0]

    def loop(a, b, stereo, proj, bow, reps):
        ts, last, counts, prev_kd = [], None, None, None
        for i in range(reps):
            l, r = pairs[i % len(pairs)]
            t, (cur, kl, dl), n, nm = one(l, r, a, b, stereo, proj, last)
            tb, nb = 0.0, 0
            if prev_kd is not None:                  # TrackReferenceKeyFrame's SearchByBoW(last as key frame, current frame)
                args = bow_inputs((kl, dl), prev_kd)
                t5 = time.perf_counter()
                nb = bow(*args)
                tb = time.perf_counter() - t5
            prev_kd = (kl, dl)
            t = t + (tb,)
            if i == 7:
                counts = (int(n), int(nm), int(nb))  # same pair and map points in both arms: the counts must agree
            pts = scenes.last_frame_points(cur, seed=i)
            last = (cur, pts, np.full(cur.N, -1, np.int32))
            ts.append(t)
        ts = np.array(ts[2:]) * 1e3
        return {"extract_pair_ms": float(np.median(ts[:, 0])), "stereo_match_ms": float(np.median(ts[:, 1])),
                "search_by_projection_ms": float(np.median(ts[:, 2])), "search_by_bow_ms": float(np.median(ts[:, 3])),
                "stereo_matches": counts[0], "projection_matches": counts[1], "bow_matches": counts[2]}

    out["b200"] = loop(exL, exR, gpu_stereo, gpu_proj, lambda *x: ORBmatcher(0.7, True, ctx).SearchByBoWFrame(*x)[0], 24)
    if with_cpu:
        from tests import oracle_lib
        oa, ob = oracle_lib.OracleExtractor(NFEAT, 1.2, 8, 20, 7), oracle_lib.OracleExtractor(NFEAT, 1.2, 8, 20, 7)
        om = oracle_lib.OracleMatcher(0.9, True)
        out["cpu_port"] = loop(oa, ob, lambda kl, dl, kr, dr: oracle_lib.oracle_stereo_matches(oa, ob, kl, dl, kr, dr, mbf, mb),
                               lambda cur, lastF, pts, obs: om.SearchByProjectionFrame(cur, lastF, pts, obs, 7.0, False)[0],
                               lambda *x: oracle_lib.OracleMatcher(0.7, True).SearchByBoWFrame(*x)[0], 8)
    pool.shutdown()
    exL.close(); exR.close(); ctx.close()
    return out


def run_vocabulary(local_rank, with_cpu):
    """Frame::ComputeBoW (ORBVocabulary::transform, levelsup 4) on the descriptors of one KITTI frame
    against a synthetic vocabulary of the ORBvoc shape (k = 10, L = 6, 10^6 words; the real file is absent), the tree descent
    alone on a device-resident batch, and MapPoint::ComputeDistinctiveDescriptors for 20 000 map points in one call."""
    import torch
    from orb_slam_b200 import ORBextractor
    from orb_slam_b200._lib import check, lib
    from orb_slam_b200.matcher import MatcherContext
    from orb_slam_b200.synth import synth_frame, synth_vocabulary
    from orb_slam_b200.vocabulary import ORBVocabulary, compute_distinctive_descriptors
    from tests import scenes
    ctx = MatcherContext(local_rank)
    arrays = synth_vocabulary(10, 6)
    voc = ORBVocabulary.from_arrays(ctx, 10, 6, *arrays)
    ex = ORBextractor(NFEAT, 1.2, 8, 20, 7, device=local_rank)
    descs = [ex(synth_frame(W_, H_, 700 + i))[1] for i in range(4)]
    ex.close()

    def med(fn, reps):
        ts = []
        for i in range(reps):
            t0 = time.perf_counter(); r = fn(i); ts.append(time.perf_counter() - t0)
        return float(np.median(ts[2:]) * 1e3), r
    out = {"vocabulary": "synthetic k=10 L=6, %d nodes, %d words (TF-IDF, L1)" % (voc.nnodes, voc.nwords)}
    t, r = med(lambda i: voc.transform(descs[i % 4], 4), 30)
    out["compute_bow_ms"] = t; out["bow_words"] = int(len(r[0])); out["featvec_nodes"] = int(len(r[2].node_ids))
    # descent only, descriptors already in HBM (a batch of 512 frames x 2000)
    big = torch.from_numpy(np.concatenate(descs)).cuda().repeat(128 * 4, 1)[: 512 * NFEAT].contiguous()
    res = torch.zeros((big.shape[0], 2), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    L = lib()
    for rep in range(6):
        if rep == 2:
            check(L.orbm_synchronize(ctx.h), "sync"); t0 = time.perf_counter()
        check(L.orbm_vocab_descend_device(ctx.h, voc._h, big.data_ptr(), big.shape[0], 4, res.data_ptr()), "descend")
    check(L.orbm_synchronize(ctx.h), "sync")
    dt = (time.perf_counter() - t0) / 4
    out["descend_batch"] = {"features": int(big.shape[0]), "ms": dt * 1e3, "features_per_s": big.shape[0] / dt,
                            "node_descriptor_GBps": big.shape[0] * 6 * 10 * 32 / dt / 1e9}
    D, start = scenes.observation_lists(20000, seed=5, max_obs=10, big=())
    t, best = med(lambda i: compute_distinctive_descriptors(ctx, D, start), 8)
    out["distinctive_descriptors"] = {"map_points": 20000, "descriptors": int(len(D)), "ms": t}
    if with_cpu:
        from tests import oracle_lib
        ov = oracle_lib.OracleVocabulary.from_arrays(10, 6, *arrays)
        t, ro = med(lambda i: ov.transform(descs[i % 4], 4), 12)
        out["cpu_port"] = {"compute_bow_ms": t, "bow_words": int(len(ro[0]))}
        t, ob = med(lambda i: oracle_lib.oracle_distinctive_descriptors(D, start), 4)
        out["cpu_port"]["distinctive_descriptors_ms"] = t
        a, b = voc.transform(descs[0], 4), ov.transform(descs[0], 4)
        out["identical_to_cpu_port"] = bool(np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and a[2].as_dict() == b[2]
                                            and np.array_equal(best, ob))
    voc.close(); ctx.close()
    return out


def run_euroc_batch(local_rank, with_cpu, nframes=1024):
    """BASELINE.json configs[2]: EuRoC-shape 752x480 monocular frame batch (1024 frames), nFeatures 1000, then
    SearchForInitialization (window 100, nnratio 0.9, rotation histogram) on every consecutive pair - extraction and
    matching both on device-resident data, no host round trip in between."""
    import torch
    from orb_slam_b200 import KEYPOINT_DTYPE, ORBextractor
    from orb_slam_b200._lib import check
    from orb_slam_b200.matcher import MatcherContext
    from orb_slam_b200.synth import synth_frame
    from tests import scenes
    w, h, nf, distinct = 752, 480, 1000, 32
    base = np.stack([synth_frame(w, h, 900, shift=(2 * i, -i)) for i in range(distinct)])
    order = [(i % (2 * distinct - 2)) for i in range(nframes)]
    order = [j if j < distinct else 2 * distinct - 2 - j for j in order]          # 0..31..1..0..: neighbours differ by one shift step
    d_img = torch.from_numpy(base).cuda()[torch.tensor(order, device="cuda")].contiguous()
    ex = ORBextractor(nf, 1.2, 8, 20, 7, device=local_rank)
    ctx = MatcherContext(local_rank)
    cap = ex.max_keypoints(w, h)
    d_k = torch.zeros((nframes, cap, 28), dtype=torch.uint8, device="cuda")
    d_d = torch.zeros((nframes, cap, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(nframes, dtype=torch.int32, device="cuda")
    d_m = torch.zeros((nframes - 1, cap), dtype=torch.int32, device="cuda"); d_nm = torch.zeros(nframes - 1, dtype=torch.int32, device="cuda")
    L = ex._L

    def extract():
        check(L.orbx_extract_batch_device(ex.handle, d_img.data_ptr(), nframes, w, h, w, w * h, d_k.data_ptr(), d_d.data_ptr(), cap,
                                          d_n.data_ptr()), "orbx_extract_batch_device")
        check(L.orbx_synchronize(ex.handle), "sync")

    def match():
        ctx.search_for_initialization_batch_device(d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(), cap, nframes, w, h,
                                                   d_matches12=d_m.data_ptr(), d_nmatches=d_nm.data_ptr())
        ctx.synchronize()
    torch.cuda.synchronize()
    te, tm = [], []
    for rep in range(5):
        t0 = time.perf_counter(); extract(); t1 = time.perf_counter(); match(); t2 = time.perf_counter()
        te.append(t1 - t0); tm.append(t2 - t1)
    te, tm = float(np.median(te[1:])), float(np.median(tm[1:]))
    nm = d_nm.cpu().numpy()
    out = {"workload": "BASELINE.json configs[2]: %d frames 752x480, nFeatures 1000; SearchForInitialization(window 100, 0.9) on %d consecutive pairs"
                       % (nframes, nframes - 1),
           "extract_ms": te * 1e3, "match_ms": tm * 1e3, "frames_per_s_extract_and_match": nframes / (te + tm),
           "pairs_per_s_match": (nframes - 1) / tm, "mean_matches_per_pair": float(nm.mean())}
    if with_cpu:
        from tests import oracle_lib
        k_h = d_k[:9].cpu().numpy().view(KEYPOINT_DTYPE).reshape(9, cap); d_h = d_d[:9].cpu().numpy(); n_h = d_n[:9].cpu().numpy()
        o = oracle_lib.OracleMatcher(0.9, True)
        ts, same = [], True
        m_h = d_m[:8].cpu().numpy()
        for p in range(8):
            F1 = scenes.make_frame(k_h[p, :n_h[p]], d_h[p, :n_h[p]], w, h); F2 = scenes.make_frame(k_h[p + 1, :n_h[p + 1]], d_h[p + 1, :n_h[p + 1]], w, h)
            prev = np.ascontiguousarray(np.stack([F1.mvKeysUn["x"], F1.mvKeysUn["y"]], 1), np.float32)
            t0 = time.perf_counter(); n_o, m_o = o.SearchForInitialization(F1, F2, prev, 100); ts.append(time.perf_counter() - t0)
            same = same and n_o == int(nm[p]) and np.array_equal(m_o, m_h[p, :n_h[p]])
        out["cpu_port"] = {"match_ms_per_pair": float(np.median(ts) * 1e3), "pairs_per_s_one_core": 1.0 / float(np.median(ts))}
        out["identical_to_cpu_port"] = bool(same)
    ex.close(); ctx.close()
    return out


def run_stereo_batch(local_rank, npairs=256):
    """BASELINE.json configs[1] as a batch: npairs KITTI-size stereo pairs; both images extracted on the device (one extractor
    per camera, as src/Frame.cc:94-103) and Frame::ComputeStereoMatches on all pairs, nothing leaving HBM in between."""
    import torch
    from orb_slam_b200 import ORBextractor
    from orb_slam_b200._lib import check
    from orb_slam_b200.synth import synth_frame
    from tests import scenes
    distinct = 8
    L_ = np.stack([synth_frame(W_, H_, 600 + i) for i in range(distinct)])
    R_ = np.stack([synth_frame(W_, H_, 600 + i, shift=(-(9 + 4 * i), 0)) for i in range(distinct)])
    rng = np.random.default_rng(1)
    R_ = np.clip(R_.astype(np.int16) + rng.integers(-3, 4, R_.shape), 0, 255).astype(np.uint8)
    idx = torch.arange(npairs, device="cuda") % distinct
    dL, dR = torch.from_numpy(L_).cuda()[idx].contiguous(), torch.from_numpy(R_).cuda()[idx].contiguous()
    exL, exR = ORBextractor(NFEAT, 1.2, 8, 20, 7, device=local_rank), ORBextractor(NFEAT, 1.2, 8, 20, 7, device=local_rank)
    cap = exL.max_keypoints(W_, H_)
    out = {}
    bufs = {}
    for name in "LR":
        bufs[name] = (torch.zeros((npairs, cap, 28), dtype=torch.uint8, device="cuda"), torch.zeros((npairs, cap, 32), dtype=torch.uint8, device="cuda"),
                      torch.zeros(npairs, dtype=torch.int32, device="cuda"))
    d_u = torch.zeros((npairs, cap), dtype=torch.float32, device="cuda"); d_z = torch.zeros((npairs, cap), dtype=torch.float32, device="cuda")
    d_nm = torch.zeros(npairs, dtype=torch.int32, device="cuda")
    mbf = float(scenes.KITTI_BF); mb = float(np.float32(scenes.KITTI_BF) / scenes.KITTI_K[0, 0])
    torch.cuda.synchronize()
    te, ts = [], []
    for rep in range(5):
        t0 = time.perf_counter()
        for ex, img, (k, d, n) in ((exL, dL, bufs["L"]), (exR, dR, bufs["R"])):
            check(ex._L.orbx_extract_batch_device(ex.handle, img.data_ptr(), npairs, W_, H_, W_, W_ * H_, k.data_ptr(), d.data_ptr(), cap,
                                                  n.data_ptr()), "orbx_extract_batch_device")
        exL.synchronize(); exR.synchronize()
        t1 = time.perf_counter()
        exL.stereo_match_batch_device(exR, npairs, bufs["L"][0].data_ptr(), bufs["L"][1].data_ptr(), bufs["L"][2].data_ptr(),
                                      bufs["R"][0].data_ptr(), bufs["R"][1].data_ptr(), bufs["R"][2].data_ptr(), cap, mbf, mb,
                                      d_u.data_ptr(), d_z.data_ptr(), d_nm.data_ptr())
        exL.synchronize()
        t2 = time.perf_counter()
        te.append(t1 - t0); ts.append(t2 - t1)
    te, ts = float(np.median(te[1:])), float(np.median(ts[1:]))
    out = {"stereo_pairs": npairs, "extract_both_ms": te * 1e3, "stereo_match_ms": ts * 1e3, "stereo_pairs_per_s": npairs / (te + ts),
           "mean_stereo_matches": float(d_nm.float().mean().item())}
    exL.close(); exR.close()
    return out


def bind_to_gpu_cpus(local_rank):
    """Multi-GPU runs: keep the rank's host thread (and the pinned buffers it first-touches) on the CPUs the driver
    reports as local to its GPU, so H2D / D2H do not cross sockets.  Best effort; a no-op when NVML has no affinity."""
    try:
        import pynvml
        pynvml.nvmlInit()
        try:
            import torch
            h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(torch.cuda.get_device_properties(local_rank).uuid)).encode())
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = {64 * i + b for i, wd in enumerate(words) for b in range(64) if (int(wd) >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
    except Exception:
        pass


def run_ours(args, rank, world, local_rank):
    import torch
    from orb_slam_b200 import KEYPOINT_DTYPE, ORBextractor
    from orb_slam_b200._lib import check

    dist = None
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", "WARN"):
            del os.environ["NCCL_DEBUG"]               # both levels print "NCCL version ..." on stdout: keep it to the one JSON line
        bind_to_gpu_cpus(local_rank)
        import torch.distributed as dist_
        dist = dist_
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    B = args.batch
    distinct = make_frames(min(64, B), rank * 64)
    reps = (B + len(distinct) - 1) // len(distinct)
    host = torch.empty((B, H_, W_), dtype=torch.uint8, pin_memory=True)
    host.numpy()[:] = np.concatenate([distinct] * reps)[:B]
    PITCH = (W_ + 127) // 128 * 128            # device-resident frames use a 128-byte aligned row pitch (HBM layout)
    d_img = torch.zeros((B, H_, PITCH), dtype=torch.uint8, device=dev)
    d_img[:, :, :W_] = host.to(dev)

    ex = ORBextractor(NFEAT, 1.2, 8, 20, 7, device=local_rank)
    L, h = ex._L, ex.handle
    cap = ex.max_keypoints(W_, H_)
    d_kps = torch.empty((B, cap, 28), dtype=torch.uint8, device=dev)
    d_desc = torch.empty((B, cap, 32), dtype=torch.uint8, device=dev)
    d_n = torch.zeros(B, dtype=torch.int32, device=dev)
    stream = torch.cuda.ExternalStream(L.orbx_stream(h), device=dev)

    def step_device():
        check(L.orbx_extract_batch_device(h, d_img.data_ptr(), B, W_, H_, PITCH, PITCH * H_, d_kps.data_ptr(),
                                          d_desc.data_ptr(), cap, d_n.data_ptr()), "orbx_extract_batch_device")

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step_device()
    check(L.orbx_synchronize(h), "orbx_synchronize")
    launches = L.orbx_launches_per_call(h)

    # ---- timed region 1: device-resident throughput (CUDA events on the launching stream)
    barrier()
    sampler = ClockSampler(local_rank)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step_device()
    e1.record(stream)
    check(L.orbx_synchronize(h), "orbx_synchronize")
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_ms = float(ms.item())
    nkp = float(d_n.float().mean().item())

    # ---- timed region 2: per-stage device times of the same steps (events inside the library)
    check(L.orbx_set_profiling(h, 1), "orbx_set_profiling")
    stage_ms = np.zeros(len(STAGES))
    t = np.zeros(len(STAGES), np.float32)
    for _ in range(args.steps):
        step_device()
        check(L.orbx_stage_times(h, t.ctypes.data, len(STAGES)), "orbx_stage_times")
        stage_ms += t
    stage_ms /= args.steps
    check(L.orbx_set_profiling(h, 0), "orbx_set_profiling")

    # ---- timed region 3: end to end through the host-buffer C ABI call (H2D + kernels + D2H)
    kps_h = torch.empty((B, cap, 28), dtype=torch.uint8, pin_memory=True)
    desc_h = torch.empty((B, cap, 32), dtype=torch.uint8, pin_memory=True)
    n_h = np.zeros(B, np.int32)
    ptrs = (C.c_void_p * B)(*[host.data_ptr() + i * W_ * H_ for i in range(B)])

    def step_e2e():
        check(L.orbx_extract_batch(h, ptrs, B, W_, H_, W_, kps_h.data_ptr(), desc_h.data_ptr(), cap, n_h.ctypes.data),
              "orbx_extract_batch")

    for _ in range(max(1, args.warmup // 2)):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    torch.cuda.synchronize()
    e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    barrier()
    clocks = sampler.stop()
    if dist is not None:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_s = float(e2e_s.item())

    # ---- what Tracking sees: one orbx_extract call per frame (pageable numpy image in, keypoints out), latency
    one = np.ascontiguousarray(host[0].numpy())
    k1 = np.empty((cap, 28), np.uint8); d1 = np.empty((cap, 32), np.uint8); n1 = C.c_int()
    lat = []
    for i in range(60):
        t0 = time.perf_counter()
        check(L.orbx_extract(h, one.ctypes.data, W_, H_, W_, k1.ctypes.data, d1.ctypes.data, cap, C.byref(n1)), "orbx_extract")
        lat.append(time.perf_counter() - t0)
    lat_ms = float(np.median(lat[10:]) * 1e3)
    # ---- the same batch call with pageable (numpy) buffers: staging through pinned memory inside the library
    pg_in = np.ascontiguousarray(host.numpy()[: min(B, 256)])
    nb = len(pg_in)
    pg_k = np.empty((nb, cap, 28), np.uint8); pg_d = np.empty((nb, cap, 32), np.uint8); pg_n = np.zeros(nb, np.int32)
    pg_ptrs = (C.c_void_p * nb)(*[pg_in.ctypes.data + i * W_ * H_ for i in range(nb)])
    for rep in range(3):
        t0 = time.perf_counter()
        check(L.orbx_extract_batch(h, pg_ptrs, nb, W_, H_, W_, pg_k.ctypes.data, pg_d.ctypes.data, cap, pg_n.ctypes.data),
              "orbx_extract_batch")
        pageable_fps = nb / (time.perf_counter() - t0)

    ham = None
    if not args.no_hamming:
        ham = run_hamming(args, rank, world, local_rank, dist)

    if rank == 0:
        frames_total = world * B * args.steps
        value = frames_total / (total_ms * 1e-3)
        balg, per_stage = algorithmic_bytes(W_, H_, nkp)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
        top = int(np.argmax(stage_ms))
        top_bytes = per_stage.get(STAGES[top], 0) * B
        achieved = top_bytes / (stage_ms[top] * 1e-3) / 1e9
        pipe_achieved = balg * B * args.steps / (total_ms * 1e-3) / 1e9 / world
        n_sel = int(n_h.sum())
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": "KITTI-stereo 1241x376 left+right (BASELINE.json configs[1]), nFeatures 2000, 8 levels, 1.2, FAST 20/7",
                       **SETTINGS, "frames_per_step_per_gpu": B, "l2": "inputs larger than L2 (batch %.0f MB)" % (B * W_ * H_ / 1e6),
                       "mean_keypoints_per_frame": nkp},
            "e2e": {"value": world * B * args.steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": B * W_ * H_,
                    "d2h_bytes_per_step": int((B + 1) * 4 + 63) // 64 * 64 + B * cap * 60, "keypoints_returned_per_step": n_sel},
            "gpu_launches": launches * args.steps,
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": STAGES[top], "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch_group": top_bytes,
                         "kernel_ms_per_step": float(stage_ms[top]),
                         "pipeline_achieved": pipe_achieved, "pipeline_frac": pipe_achieved / peak,
                         "algorithmic_bytes_per_frame": balg},
            "stages_ms_per_step": {k: float(v) for k, v in zip(STAGES, stage_ms)},
            "single_frame_latency_ms": lat_ms,
            "e2e_pageable_frames_per_s": pageable_fps,
        }
        line["roofline"]["traffic"] = measured_traffic(B)
        line["roofline"]["alu_bound_note"] = ("K1 is ALU-pipe bound, not HBM bound (about 120 integer instructions per pixel against 4 B "
                                              "of traffic): see ncu_pipe for its pipe utilisation")
        line["roofline"]["ncu_pipe"] = ncu_pipe_utilisation()
        if ham is not None:
            line["hamming_2nn"] = ham
        if world == 1 and not args.no_frontend:
            line["kitti_stereo_frontend"] = run_stereo_frontend(local_rank, not args.no_cpu)
            line["kitti_stereo_frontend"]["batch"] = run_stereo_batch(local_rank)
            line["vocabulary"] = run_vocabulary(local_rank, not args.no_cpu)
            line["euroc_batch"] = run_euroc_batch(local_rank, not args.no_cpu)
        if world == 1 and not args.no_cpu:
            from tests import oracle_lib
            kind = "reference" if oracle_lib.have_ref() else "port"
            cores = os.cpu_count() or 1
            per_thread = max(2, args.cpu_frames // cores)
            v, dt = cpu_extract_throughput(distinct, cores, per_thread, kind)
            v1, _ = cpu_extract_throughput(distinct, 1, 8, kind)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": kind,
                                    "sample": f"{cores} threads x {per_thread} frames of the same workload ({dt:.1f} s wall)",
                                    "single_thread_frames_per_s": v1}
            if ham is not None:
                line["hamming_2nn"]["cpu_baseline"] = cpu_hamming_baseline(cores)
        print(json.dumps(line), flush=True)
    ex.close()
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=512, help="frames per step per GPU")
    ap.add_argument("--cpu-frames", type=int, default=768, help="frames in the bounded CPU sample")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-hamming", action="store_true")
    ap.add_argument("--no-frontend", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
