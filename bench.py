#!/usr/bin/env python
"""bench.py — ORB front-end throughput on B200 (BASELINE.json metric: ORB frames/s + Hamming kNN Gpairs/s; %roofline).

One "step" = one pass of the hot path over one batch of synthetic input.  Default workload = BASELINE.json
configs[1] ("c2"): EuRoC-shaped 752x480 stereo pairs, 1200 features per image, 512 pairs per GPU: extraction of the
1024 images (pyramid, per-cell FAST, quadtree, orientation, blur, rBRIEF) + left->right brute-force Hamming kNN (k=2).
`value` is measured with the inputs resident in HBM; `e2e` goes through the same C-ABI calls with pinned HOST buffers
(host->device and device->host copies inside the timed region).  `--impl reference` times the CPU oracle
(oracle/cpp, the restatement of the reference's ORBextractor_old.cc + OpenCV primitives, pinned bit for bit against the
reference's own compiled extractor oracle/_ref/libext_old.so) on the host cores on a bounded sample of the
same workload; the reference build itself is timed beside it single-threaded (`cpu_baseline.c1_single_frame_1thread`).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H, NFEAT, NLEVELS = 752, 480, 1200, 8
# algorithmic bytes per 752x480 image (DESIGN.md "Roofline")
ALGO_BYTES = {"pyramid": 1845634, "fast": 1117367, "blur": 2234734}


def make_pairs(npairs, seed0):
    from orbslam3lib_b200 import scenes
    imgs = np.empty((2 * npairs, H, W), np.uint8)
    for k in range(npairs):
        left = scenes.make("texture", W, H, seed0 + k)
        imgs[k] = left
        imgs[npairs + k] = scenes.stereo_right(left, seed0 + k)
    return imgs


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


# kNN roofline: tensor-bound, 512 int8 operations per descriptor pair.  Peak = 2 x the MEASURED dense bf16 rate of
# MEASURED_PEAKS.json (tcgen05 kind::i8 issues K = 32 bytes per instruction against K = 16 bf16 elements for kind::f16:
# twice the operations per instruction at the same instruction rate); fallback 2 x 2250 nominal.  The legacy mma.sync
# IMMA issue peak measured with tools/ubench/mma_probe.cu (1133 Tops/s) is the ceiling of the ORBFE_KNN_TC=0 kernel.
IMMA_SYNC_PEAK_TOPS = 1133.0
MAP_1GPU_MS = 179.8    # 1M x 1M kNN on one B200 (SCALE_r01.json, N=1): the denominator of knn_map_summary.speedup


def int8_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        if "bf16_tflops" in d:
            return (2.0 * float(d["bf16_tflops"]), 2.0 * float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                    "measured: 2 x MEASURED_PEAKS.json bf16_tflops (burst); sustained = 2 x bf16_tflops_sustained")
    return 4500.0, 4500.0, "fallback: 2 x 2250 nominal dense bf16"


def knn_roofline(gpairs, world):
    tops = gpairs * 1e9 * 512 * 1e-12
    burst, sustained, src = int8_peak()
    return {"bound": "tensor", "achieved": round(tops, 1), "peak": burst * world, "unit": "Tops/s (int8, tcgen05 kind::i8)",
            "frac": round(tops / (burst * world), 4), "peak_sustained": sustained * world,
            "frac_of_sustained": round(tops / (sustained * world), 4), "peak_source": src,
            "mma_sync_imma_peak": IMMA_SYNC_PEAK_TOPS * world, "popc_kernel_ceiling_gpairs": round(world * 148 * 2.0 * 1.92, 1)}


class ClockSampler(threading.Thread):
    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        self.proc = None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                if self.stop_flag:
                    break
                f = [x.strip() for x in line.split(",")]
                if len(f) < 6:
                    continue
                try:
                    self.samples.append(float(f[0]))
                    self.max_mhz = float(f[1])
                except ValueError:
                    continue
                for n, v in zip(names, f[2:6]):
                    if v.lower().startswith("active"):
                        self.reasons.add(n)
        except Exception:
            pass

    def stop(self):
        self.stop_flag = True
        if self.proc:
            self.proc.terminate()
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def cpu_sample(npairs_sample, threads, budget_s=12.0):
    """The CPU oracle on a bounded sample of the workload: frames/s and Gpairs/s on `threads` host threads."""
    sys.path.insert(0, os.path.join(ROOT, "oracle", "py"))
    import oracle_lib
    oracle_lib.build()
    imgs = make_pairs(npairs_sample, 1000)
    t0 = time.perf_counter()
    frames, pairs, reps = 0, 0, 0
    while True:
        counts, monos, kps, desc = oracle_lib.extract_batch(imgs, NFEAT, 1.2, NLEVELS, 20, 7, threads=threads)
        for k in range(npairs_sample):
            nq, nt = int(counts[k]), int(counts[npairs_sample + k])
            oracle_lib.knn2_hamming(desc[k, :nq], desc[npairs_sample + k, :nt], threads=threads)
            pairs += nq * nt
        frames += 2 * npairs_sample
        reps += 1
        if time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    return frames / dt, pairs / dt * 1e-9, "%d stereo pairs 752x480 nf=1200 x %d repetitions (%.1f s)" % (npairs_sample, reps, dt)


def cpu_extra(cores):
    """The CPU legs BASELINE.md §2 / north_star name beside the multi-threaded C2 number (bounded: ~15 s in total):
    C1 single frame on one thread with the per-stage split (the log points of ORBextractor_old.cc:1107,1113,1149,1156,
    1187 = the oracle's stage timers) and the same frame through the reference's own compiled extractor (oracle/_ref)
    when that library travelled with the snapshot; C3 stereo pair on 2 threads, p50 / p99; the matcher on 1 thread and on
    all cores at the C2 shape and on a 4096-query sample of the C5 shape; a cv2.BFMatcher cross-check where cv2 imports."""
    import oracle_lib
    from orbslam3lib_b200 import scenes
    out = {}
    # C1: 640x480, nf=1000, 8 levels, one thread
    img = scenes.make("texture", 640, 480, 1000)
    ex = oracle_lib.OracleExtractor(1000, 1.2, 8, 20, 7)
    ex(img)
    s0 = ex.stage_seconds()
    t0 = time.perf_counter()
    reps = 10
    for _ in range(reps):
        _, k1, d1 = ex(img)
    dt = (time.perf_counter() - t0) / reps
    s1 = ex.stage_seconds()
    out["c1_single_frame_1thread"] = {
        "workload": "c1: 640x480, nf=1000, 8 levels, ORBextractor::operator() on one thread (CPU oracle)",
        "ms_per_frame": dt * 1e3, "keypoints": int(len(k1)),
        "stage_ms": {k: (s1[k] - s0[k]) / reps * 1e3 for k in s1}}
    try:
        import ref_ext_lib
        if ref_ext_lib.available():
            rex = ref_ext_lib.RefExtractor(1000, 1.2, 8, 20, 7)
            rex(img)
            t0 = time.perf_counter()
            for _ in range(reps):
                _, rk, rd = rex(img)
            out["c1_single_frame_1thread"]["reference_build_ms_per_frame"] = (time.perf_counter() - t0) / reps * 1e3
            out["c1_single_frame_1thread"]["reference_build"] = ("oracle/_ref/libext_old.so = the reference's own "
                                                                 "ORBextractor_old.cc (g++ -O2) over the oracle's primitives")
            out["c1_single_frame_1thread"]["reference_build_equals_oracle"] = bool(
                np.array_equal(rk, k1) and np.array_equal(rd, d1))
    except Exception as e:    # the reference build is optional on the GPU box
        out["c1_single_frame_1thread"]["reference_build"] = "unavailable: %s" % e
    # matcher against the previous frame (configs[0]): k=2 on one thread
    _, k0, d0 = ex(scenes.previous_frame(img, 1000))
    t0 = time.perf_counter()
    for _ in range(20):
        oracle_lib.knn2_hamming(d1, d0, threads=1)
    out["c1_single_frame_1thread"]["knn_prev_frame_ms"] = (time.perf_counter() - t0) / 20 * 1e3
    # C3: 640x480 stereo pair, nf=500, two threads (one per eye, as Frame.cc:142-145 does)
    left = scenes.make("texture", 640, 480, 3000)
    pair = np.stack([left, scenes.stereo_right(left, 3000)])
    ts = []
    for it in range(40):
        t0 = time.perf_counter()
        counts, _, _, desc = oracle_lib.extract_batch(pair, 500, 1.2, 8, 20, 7, threads=2)
        oracle_lib.knn2_hamming(desc[0, :counts[0]], desc[1, :counts[1]], threads=1)
        if it >= 5:
            ts.append((time.perf_counter() - t0) * 1e3)
    ts.sort()
    out["c3_stereo_pair_2threads"] = {"workload": "c3: 640x480 stereo pair, nf=500, extraction on 2 threads + L->R kNN",
                                      "p50_ms": ts[len(ts) // 2], "p99_ms": ts[-1], "iters": len(ts)}
    # matcher: C2 shape (1200 x 1200) on one thread and on all cores; C5 shape on a 4096-query sample
    q, t = scenes.map_descriptors(1200, 1200, 9100)
    knn = {}
    for name, th in (("1thread", 1), ("%dthreads" % cores, cores)):
        t0 = time.perf_counter()
        reps = 20 if th == 1 else 100
        for _ in range(reps):
            oracle_lib.knn2_hamming(q, t, threads=th)
        knn["c2_shape_gpairs_per_s_" + name] = 1200 * 1200 * reps / (time.perf_counter() - t0) * 1e-9
    q5, t5 = scenes.map_descriptors(4096, 1000000, 9000)
    t0 = time.perf_counter()
    i5, d5 = oracle_lib.knn2_hamming(q5, t5, threads=cores)
    dt5 = time.perf_counter() - t0
    knn["c5_sample"] = {"workload": "4096 query rows x 1M train rows (a sample of configs[4]), %d threads" % cores,
                        "seconds": dt5, "gpairs_per_s": 4096 * 1e6 / dt5 * 1e-9}
    try:
        import cv2
        # cv2.BFMatcher refuses train sets of 2^18 rows or more (IMGIDX_ONE): cross-check on the first 262143 rows
        nt = 262143
        m = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(q5[:64], t5[:nt], k=2)
        ci = np.array([[d.trainIdx for d in row] for row in m], np.int32)
        cd = np.array([[int(d.distance) for d in row] for row in m], np.int32)
        oi, od = oracle_lib.knn2_hamming(q5[:64], t5[:nt], threads=cores)
        knn["cv2_bfmatcher_crosscheck"] = {"rows": 64, "train_rows": nt, "cv2": cv2.__version__,
                                           "equal": bool(np.array_equal(ci, oi) and np.array_equal(cd, od))}
    except ImportError:
        knn["cv2_bfmatcher_crosscheck"] = "cv2 not importable on this box (checked in the build container: tests/test_oracle_vs_cv2.py)"
    out["knn"] = knn
    return out


def map_knn_leg(args, rank, world, local, dev, dist, matcher):
    """BASELINE configs[4]: brute-force Hamming kNN k=2 of nq x nt 256-bit descriptors, query rows sharded over the
    ranks, train set broadcast and results exchanged with NCCL (orbfe_knn2_sharded).  Strong scaling: the problem is
    fixed, so the per-step time falls with N.  Device time, max over ranks."""
    import torch
    from orbslam3lib_b200 import ShardedMatcher, scenes, shard_range
    nq = nt = int(args.map_rows)
    q, t = scenes.map_descriptors(nq, nt, 9000)    # every rank derives the same seeded sets; only the root uploads t
    b, c = shard_range(nq, world, rank)
    d_q = torch.from_numpy(q[b:b + c].copy()).to(dev)
    d_t = torch.from_numpy(t).to(dev) if rank == 0 else torch.zeros((nt, 32), dtype=torch.uint8, device=dev)
    d_idx = torch.full((nq, 2), -9, dtype=torch.int32, device=dev)
    d_dist = torch.full((nq, 2), -9, dtype=torch.int32, device=dev)
    sm = ShardedMatcher(matcher, rank, world, local, dist)
    times, parts = [], {"kernel_ms": 0.0, "bcast_ms": 0.0, "gather_ms": 0.0}
    for it in range(1 + args.map_steps):
        if world > 1:
            if rank != 0:
                d_t.zero_()    # the broadcast has to deliver the train set every step
            dist.barrier()
        torch.cuda.synchronize()
        ms = sm.knn2(d_q, nq, d_t, nt, d_idx, d_dist)
        if world > 1:
            tt = torch.tensor([ms], device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms = float(tt.item())
        if it > 0:
            times.append(ms)
            for k, v in sm.last_times().items():
                if k in parts:
                    parts[k] += v / args.map_steps    # this rank's split of the step
    idx = d_idx.cpu().numpy()
    dd = d_dist.cpu().numpy()
    out = None
    if rank == 0:
        sys.path.insert(0, os.path.join(ROOT, "oracle", "py"))
        import oracle_lib
        sel = np.arange(0, nq, max(1, nq // 256))[:256]
        ri, rd = oracle_lib.knn2_hamming(q[sel], t, threads=oracle_lib.hardware_threads())
        ok = bool(np.array_equal(idx[sel], ri) and np.array_equal(dd[sel], rd))
        ms = sum(times) / len(times)
        _, _, sm_mhz = peaks()
        gp = nq * nt / (ms * 1e-3) * 1e-9
        out = {"workload": "c5: %d x %d 256-bit descriptors, k=2, query rows sharded over %d GPU(s), NCCL train "
                           "broadcast + result exchange" % (nq, nt, world),
               "ms_per_step": ms, "gpairs_per_s": gp, "scaling": "strong", "steps": args.map_steps,
               "kernel_ms": round(parts["kernel_ms"], 3), "comm_ms": round(ms - parts["kernel_ms"], 3),
               "bcast_ms_own_stream": round(parts["bcast_ms"], 3), "gather_ms": round(parts["gather_ms"], 3),
               "split_note": "rank 0: kernel_ms = the kNN launches, comm_ms = max-over-ranks step - kernel_ms (exposed "
                             "communication + waiting for the slowest rank); the train broadcast runs on its own stream "
                             "under the kernels",
               "roofline": knn_roofline(gp, world),
               "checksum": int((idx.astype(np.int64) * 31 + dd).sum() & 0xFFFFFFFFFFFF),
               "parity_sample": {"rows": int(len(sel)), "vs": "CPU oracle", "ok": ok}}
    sm.close()
    del d_q, d_t, d_idx, d_dist
    torch.cuda.empty_cache()
    return out


def latency_leg(args, local):
    """BASELINE configs[2]: Lynx-R1-style 640x480 stereo stream, 500 features per image: per-frame latency of one
    stereo pair through the C ABI with HOST buffers (copy-in, extraction of both eyes, left->right kNN, copy-out),
    wall clock around the blocking calls."""
    import torch
    from orbslam3lib_b200 import BFMatcher, ORBextractor, _lib, scenes
    w, h, nf = 640, 480, 500
    frames = []
    for k in range(8):
        left = scenes.make("texture", w, h, 3000 + k)
        frames.append(torch.from_numpy(np.stack([left, scenes.stereo_right(left, 3000 + k)])).pin_memory())
    ex = ORBextractor(nf, 1.2, 8, 20, 7, device=local)
    mt = BFMatcher(device=local)
    cap = ex.max_keypoints
    kps = torch.zeros((2, cap, 28), dtype=torch.uint8).pin_memory()
    desc = torch.zeros((2, cap, 32), dtype=torch.uint8).pin_memory()
    counts = torch.zeros(2, dtype=torch.int32).pin_memory()
    monos = torch.zeros(2, dtype=torch.int32).pin_memory()
    idx = torch.zeros((cap, 2), dtype=torch.int32).pin_memory()
    dist = torch.zeros((cap, 2), dtype=torch.int32).pin_memory()
    ts = []
    for it in range(args.latency_iters + 20):
        f = frames[it % len(frames)]
        t0 = time.perf_counter()
        # one call per pair: both eyes + left -> right kNN; pinned buffers, so the call replays a CUDA graph from its
        # second use of a buffer slot on (csrc/orbfe_api.cu, ORBFE_GRAPH=0 issues the ~18 stream operations one by one)
        ex.extract_match_batch_raw(f, 1, w, h, kps, desc, cap, counts, monos, idx, dist)
        t1 = time.perf_counter()
        if it >= 20:
            ts.append((t1 - t0) * 1e3)
    ts.sort()
    return {"workload": "c3: 640x480 stereo pair, 500 features/image, pinned host buffers, one orbfe_extract_match_batch call "
                        "per pair (extraction of both eyes + L->R kNN), CUDA-graph replay",
            "cuda_graph": os.environ.get("ORBFE_GRAPH", "1") != "0",
            "iters": len(ts), "p50_ms": ts[len(ts) // 2], "p99_ms": ts[min(len(ts) - 1, int(len(ts) * 0.99))],
            "mean_ms": sum(ts) / len(ts), "fps_at_p50": 1e3 / ts[len(ts) // 2],
            "keypoints": [int(counts[0]), int(counts[1])]}


def backend_leg(args, local, dev):
    """SURVEY.md 8(f) rank 3, the matchers of the local-mapping / loop-closing threads: the Sim3 projection
    search, Fuse, SearchBySim3, SearchByBoW (key frame -> frame and key frame -> key frame), SearchForTriangulation and
    MapPoint::ComputeDistinctiveDescriptors, one batch of args.search_frames key frames (pairs) per step, everything
    resident in HBM; the CPU oracle is timed beside every one on one thread and checked on the sampled key frames."""
    import torch
    from orbslam3lib_b200 import ORBextractor, ORBmatcher, _lib, scenes
    sys.path.insert(0, os.path.join(ROOT, "oracle", "py"))
    import oracle_lib
    oracle_lib.build()
    nfr, npts, steps = args.search_frames, 2000, 5
    ex = ORBextractor(NFEAT, 1.2, NLEVELS, 20, 7, device=local)
    cap = ex.max_keypoints
    nsc = min(nfr, 8)
    cur = np.stack([scenes.make("texture", W, H, 7400 + k) for k in range(nsc)])
    prev = np.stack([scenes.previous_frame(cur[k], 7400 + k) for k in range(nsc)])
    c1, _, k1s, d1s = ex.extract_batch(cur)
    c2, _, k2s, d2s = ex.extract_batch(prev)
    scale = np.asarray(ex.GetScaleFactors(), np.float32)
    sigma2 = (scale * scale).astype(np.float32)
    inv_sigma2 = (1.0 / sigma2).astype(np.float32)
    bounds = (0.0, 0.0, float(W), float(H))
    FL = _lib.IN_DEVICE | _lib.OUT_DEVICE
    S = []    # per sampled scene: everything the oracle needs
    for k in range(nsc):
        a, b = int(c1[k]), int(c2[k])
        ka, da, kb, db = k1s[k, :a], d1s[k, :a], k2s[k, :b], d2s[k, :b]
        r1, r2 = scenes.rand_u32(a, 7500 + k), scenes.rand_u32(b, 7600 + k)
        S.append(dict(k1=ka, d1=da, k2=kb, d2=db, u1=np.where(r1 % 3 == 0, -1.0, ka["x"] - 8.0).astype(np.float32),
                      has1=(r1 % 5 != 0).astype(np.uint8), has2=(r2 % 5 != 0).astype(np.uint8),
                      fv1=scenes.feature_vector(da), fv2=scenes.feature_vector(db),
                      pts=scenes.frame_points(ka, da, a, None, None, NLEVELS, npts, 7700 + k, dup_frac=0.2),
                      p1=scenes.frame_points(kb, db, b, None, None, NLEVELS, a, 7800 + k, dup_frac=0.05),
                      p2=scenes.frame_points(ka, da, a, None, None, NLEVELS, b, 7900 + k, dup_frac=0.05)))
        c, rm = S[-1], scenes.rand_u32(a, 8000 + k)
        for i1 in range(0, a, 2):    # mutual pairs: MapPoint i1 of KF1 projects onto feature j of KF2 and j's projects back
            j = int(rm[i1] % b)
            for pp, i, kk, dd, o in ((c["p1"], i1, kb, db, j), (c["p2"], j, ka, da, i1)):
                pp[i]["u"], pp[i]["v"], pp[i]["level"], pp[i]["desc"], pp[i]["flags"] = kk["x"][o], kk["y"][o], kk["octave"][o], dd[o], 1
    ncap = max(max(len(c["fv1"][0]), len(c["fv2"][0])) for c in S)
    spec = dict(k=((nfr, cap), _lib.KP_DTYPE), d=((nfr, cap, 32), np.uint8), a=((nfr, cap), np.float32),
                v=((nfr, cap), np.uint8), fr=((nfr, cap), np.uint8), id=((nfr, ncap), np.int32),
                st=((nfr, ncap + 1), np.int32), ix=((nfr, cap), np.int32), nn=((nfr,), np.int32), n=((nfr,), np.int32),
                p=((nfr, cap), scenes.FRAME_POINT_DTYPE))
    A = {side: {key: np.zeros(shp, dt) for key, (shp, dt) in spec.items()} for side in ("1", "2")}
    U = np.full((nfr, cap), -1, np.float32)
    P = np.zeros((nfr, npts), scenes.FRAME_POINT_DTYPE)
    for f in range(nfr):
        c = S[f % nsc]
        for side in ("1", "2"):
            kk, dd, fv, X = c["k" + side], c["d" + side], c["fv" + side], A[side]
            n = len(kk)
            X["k"][f, :n], X["d"][f, :n], X["a"][f, :n], X["v"][f, :n], X["n"][f] = kk, dd, kk["angle"], c["has" + side], n
            X["fr"][f, :n] = 1 - c["has" + side]
            X["id"][f, :len(fv[0])], X["st"][f, :len(fv[1])], X["ix"][f, :len(fv[2])], X["nn"][f] = fv[0], fv[1], fv[2], len(fv[0])
            X["p"][f, :n] = c["p" + side]
        U[f, :len(c["k1"])] = c["u1"]
        P[f] = c["pts"]
    tt = lambda a: torch.from_numpy(a.view(np.uint8).reshape(a.shape[0], -1) if a.dtype.fields else a).to(dev)    # noqa: E731
    T = {side: {key: tt(v) for key, v in A[side].items()} for side in ("1", "2")}
    a, b = T["1"], T["2"]
    dU, dP = torch.from_numpy(U).to(dev), tt(P)
    dnp = torch.full((nfr,), npts, dtype=torch.int32, device=dev)
    z = lambda *shp: torch.zeros(shp, dtype=torch.int32, device=dev)    # noqa: E731
    out = {}

    def timed(call):
        ms = 0.0
        for it in range(steps + 2):
            mobj = call()
            if it >= 2:
                ms += mobj.last_ms()[0]
        return ms / steps

    def cpu(fn):
        t0 = time.perf_counter()
        res = [fn(c) for c in S]
        return res, nsc / (time.perf_counter() - t0)

    def entry(name, workload, ms, got, ref, cpu_rate, extra=None):
        ok = all(bool(np.array_equal(g, r)) for g, r in zip(got, ref))
        out[name] = {"workload": workload, "kernel_ms_per_step": ms, "key_frames_per_s": nfr / (ms * 1e-3),
                     "cpu_oracle_key_frames_per_s_1thread": cpu_rate,
                     "parity_sample": {"key_frames": nsc, "vs": "CPU oracle (bit-exact indices)", "ok": ok}}
        if extra:
            out[name].update(extra)

    m75, m60 = ORBmatcher(0.75, True, device=local), ORBmatcher(0.6, True, device=local)
    # SearchByProjection(pKF, Scw, vpPoints, vpMatched, th = 10, ratioHamming = 1) (ORBmatcher.cc:428-647)
    d_m, d_own, d_cnt = z(nfr, npts), z(nfr, cap), z(nfr)

    def run_sim3p():
        m75.search_sim3_projection_batch_raw(a["k"], a["d"], None, a["n"], cap, bounds, scale, dP, dnp, npts, nfr, 10.0, 1.0,
                                             d_m, d_own, d_cnt, FL)
        return m75._bf
    ms = timed(run_sim3p)
    got = d_m[:nsc].cpu().numpy()
    ref, rate = cpu(lambda c: oracle_lib.search_by_projection_sim3(c["k1"], c["d1"], None, bounds, scale, c["pts"], 10.0, 1.0)[0])
    entry("sim3_projection", "ORBmatcher::SearchByProjection(pKF, Scw, ...), %d key frames x %d points, th = 10" % (nfr, npts),
          ms, got, ref, rate, {"matches_per_key_frame": float(d_cnt.float().mean().item())})
    # Fuse(pKF, vpMapPoints, th = 3) (ORBmatcher.cc:1155-1345)
    d_bi, d_bd = z(nfr, npts), z(nfr, npts)

    def run_fuse():
        m60.fuse_batch_raw(a["k"], a["d"], dU, a["n"], cap, bounds, scale, inv_sigma2, 47.9, dP, dnp, npts, nfr, 3.0, True, d_bi,
                           d_bd, d_cnt, FL)
        return m60._bf
    ms = timed(run_fuse)
    got = d_bi[:nsc].cpu().numpy()
    ref, rate = cpu(lambda c: oracle_lib.fuse(c["k1"], c["d1"], c["u1"], bounds, scale, inv_sigma2, 47.9, c["pts"], 3.0, True)[0])
    entry("fuse", "ORBmatcher::Fuse(pKF, vpMapPoints, th = 3), %d key frames x %d points, chi-square gate" % (nfr, npts), ms,
          got, ref, rate, {"candidates_per_key_frame": float(d_cnt.float().mean().item())})
    return backend_leg_pairs(args, local, dev, S, a, b, cap, ncap, nfr, nsc, bounds, scale, sigma2, timed, cpu, entry, out, m75,
                             m60, z, FL)


def backend_leg_pairs(args, local, dev, S, a, b, cap, ncap, nfr, nsc, bounds, scale, sigma2, timed, cpu, entry, out, m75, m60,
                      z, FL):
    """second half of backend_leg: the key-frame pair matchers and the distinctive-descriptor selection."""
    import torch
    from orbslam3lib_b200 import scenes
    import oracle_lib
    d_m12, d_v1, d_v2, d_cnt = z(nfr, cap), z(nfr, cap), z(nfr, cap), z(nfr)
    # SearchBySim3(pKF1, pKF2, vpMatches12, S12, th = 7.5) (ORBmatcher.cc:1464-1681)

    def run_sim3():
        m75.search_by_sim3_batch_raw(a["k"], a["d"], a["n"], b["k"], b["d"], b["n"], cap, bounds, bounds, scale, a["p"], b["p"],
                                     nfr, 7.5, d_m12, d_v1, d_v2, d_cnt, FL)
        return m75
    ms = timed(run_sim3)
    g = d_m12[:nsc].cpu().numpy()
    got = [g[k, :len(S[k]["k1"])] for k in range(nsc)]
    ref, rate = cpu(lambda c: oracle_lib.search_by_sim3(c["k1"], c["d1"], c["k2"], c["d2"], bounds, bounds, scale, c["p1"], c["p2"],
                                                        7.5)[0])
    entry("search_by_sim3", "ORBmatcher::SearchBySim3, %d key-frame pairs, ~%d MapPoints each way, th = 7.5"
          % (nfr, len(S[0]["k1"])), ms, got, ref, rate, {"found_per_pair": float(d_cnt.float().mean().item())})
    # SearchByBoW(pKF, F, vpMapPointMatches) (ORBmatcher.cc:224-425): key frame = side 2 (previous frame), frame = side 1

    def run_bow():
        m75.search_by_bow_batch_raw(b["d"], b["a"], b["v"], b["id"], b["st"], b["ix"], b["nn"], a["d"], a["a"], a["n"], None,
                                    a["id"], a["st"], a["ix"], a["nn"], cap, ncap, nfr, d_m12, d_cnt, FL)
        return m75
    ms = timed(run_bow)
    g = d_m12[:nsc].cpu().numpy()
    got = [g[k, :len(S[k]["k1"])] for k in range(nsc)]
    ref, rate = cpu(lambda c: oracle_lib.search_by_bow(c["d2"], c["k2"]["angle"].astype(np.float32), c["has2"], c["fv2"], c["d1"],
                                                       c["k1"]["angle"].astype(np.float32), -1, c["fv1"], 0.75, True)[0])
    entry("search_by_bow", "ORBmatcher::SearchByBoW(KeyFrame, Frame), %d pairs, %d vocabulary nodes" % (nfr, ncap), ms, got,
          ref, rate, {"matches_per_pair": float(d_cnt.float().mean().item())})
    # SearchByBoW(pKF1, pKF2, vpMatches12) (ORBmatcher.cc:766-906)

    def run_bow_kf():
        m75.search_by_bow_kf_batch_raw(a["d"], a["a"], a["v"], a["id"], a["st"], a["ix"], a["nn"], b["d"], b["a"], b["v"], b["id"],
                                       b["st"], b["ix"], b["nn"], cap, ncap, nfr, d_m12, d_cnt, FL)
        return m75
    ms = timed(run_bow_kf)
    g = d_m12[:nsc].cpu().numpy()
    got = [g[k, :len(S[k]["k1"])] for k in range(nsc)]
    ang = lambda c, s_: c["k" + s_]["angle"].astype(np.float32)    # noqa: E731
    ref, rate = cpu(lambda c: oracle_lib.search_by_bow_kf(c["d1"], ang(c, "1"), c["has1"], c["fv1"], c["d2"], ang(c, "2"),
                                                          c["has2"], c["fv2"], 0.75, True)[0])
    entry("search_by_bow_kf", "ORBmatcher::SearchByBoW(KeyFrame, KeyFrame), %d pairs" % nfr, ms, got, ref, rate,
          {"matches_per_pair": float(d_cnt.float().mean().item())})
    # SearchForTriangulation (ORBmatcher.cc:908-1153): pure image translation (3, 2) between the key frames
    F12 = np.tile(np.array([0, 0, 2, 0, 0, -3, -2, 3, 0], np.float32), (nfr, 1))
    EP = np.tile(np.array([W * 0.4, H * 0.45], np.float32), (nfr, 1))
    dF, dE = torch.from_numpy(F12).to(dev), torch.from_numpy(EP).to(dev)

    def run_tri():
        m60.search_for_triangulation_batch_raw(a["k"], a["d"], a["fr"], None, a["id"], a["st"], a["ix"], a["nn"], b["k"], b["d"],
                                               b["fr"], None, b["id"], b["st"], b["ix"], b["nn"], cap, ncap, nfr, dF, dE, scale,
                                               sigma2, False, d_m12, d_cnt, FL)
        return m60
    ms = timed(run_tri)
    g = d_m12[:nsc].cpu().numpy()
    got = [g[k, :len(S[k]["k1"])] for k in range(nsc)]
    ref, rate = cpu(lambda c: oracle_lib.search_for_triangulation(c["k1"], c["d1"], 1 - c["has1"], None, c["fv1"], c["k2"], c["d2"],
                                                                  1 - c["has2"], None, c["fv2"], F12[0], EP[0], scale, sigma2,
                                                                  False, True)[0])
    entry("search_for_triangulation", "ORBmatcher::SearchForTriangulation, %d pairs, pinhole epipolar constraint" % nfr, ms, got,
          ref, rate, {"matches_per_pair": float(d_cnt.float().mean().item())})
    # MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:331-401): 50k MapPoints with 2 .. 17 observations each
    npt = 50000
    cnts = (scenes.rand_u32(npt, 8100) % np.uint32(16)).astype(np.int64) + 2
    off = np.concatenate([[0], np.cumsum(cnts)]).astype(np.int32)
    rows, _ = scenes.map_descriptors(int(off[-1]), 4096, 8101)
    d_rows, d_off = torch.from_numpy(rows).to(dev), torch.from_numpy(off).to(dev)
    d_bi, d_bm = z(npt), z(npt)
    d_od = torch.zeros((npt, 32), dtype=torch.uint8, device=dev)
    from orbslam3lib_b200 import _lib

    def run_dd():
        m60._bf._check(_lib.lib().orbfe_distinctive_descriptors(m60._bf._h, _lib.ptr(d_rows), _lib.ptr(d_off), npt, len(rows), 32,
                                                                _lib.ptr(d_bi), _lib.ptr(d_bm), _lib.ptr(d_od), FL))
        return m60
    ms = timed(run_dd)
    nsam = 2000
    t0 = time.perf_counter()
    ri, rm = oracle_lib.distinctive_descriptors(rows[:off[nsam]], off[:nsam + 1])
    t_cpu = time.perf_counter() - t0
    ok = bool(np.array_equal(d_bi[:nsam].cpu().numpy(), ri) and np.array_equal(d_bm[:nsam].cpu().numpy(), rm))
    out["distinctive_descriptors"] = {"workload": "MapPoint::ComputeDistinctiveDescriptors, %d MapPoints, %d observations"
                                                  % (npt, len(rows)), "kernel_ms_per_step": ms,
                                      "map_points_per_s": npt / (ms * 1e-3), "cpu_oracle_map_points_per_s_1thread": nsam / t_cpu,
                                      "parity_sample": {"map_points": nsam, "vs": "CPU oracle (BestIdx, BestMedian)", "ok": ok}}
    return out


def stereo_leg(args, local, dev):
    """the rectified-stereo front-end of Frame::Frame (two extractions) followed by
    Frame::ComputeStereoMatches (Frame.cc:827-997) on device-resident pyramids, P pairs per step
    (orbfe_extract_stereo_batch, images interleaved L, R).  Device time of the two stereo kernels from CUDA events
    inside the library, whole call from torch events; the CPU oracle's orc_stereo_matches timed beside it."""
    import torch
    from orbslam3lib_b200 import ORBextractor, _lib, scenes
    P = args.stereo_pairs
    mb, mbf = 0.11, 0.11 * 435.2
    imgs = np.empty((P, 2, H, W), np.uint8)
    nsc = min(P, 16)
    for k in range(nsc):
        imgs[k, 0] = scenes.make("texture", W, H, 5000 + k)
        imgs[k, 1] = scenes.stereo_right_rectified(imgs[k, 0], 5000 + k)
    for k in range(nsc, P):
        imgs[k] = imgs[k % nsc]
    d_imgs = torch.from_numpy(imgs).to(dev)
    ex = ORBextractor(NFEAT, 1.2, NLEVELS, 20, 7, device=local)
    ex.set_stream(torch.cuda.current_stream().cuda_stream)
    cap = ex.max_keypoints
    d_kps = torch.zeros((2 * P, cap, 28), dtype=torch.uint8, device=dev)
    d_desc = torch.zeros((2 * P, cap, 32), dtype=torch.uint8, device=dev)
    d_counts = torch.zeros(2 * P, dtype=torch.int32, device=dev)
    d_u = torch.zeros((P, cap), dtype=torch.float32, device=dev)
    d_z = torch.zeros((P, cap), dtype=torch.float32, device=dev)
    FL = _lib.IN_DEVICE | _lib.OUT_DEVICE
    ms_st, ms_all, steps = 0.0, 0.0, 3
    for it in range(steps + 2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        ex.extract_stereo_batch_raw(d_imgs, P, W, H, d_kps, d_desc, cap, d_counts, mb, mbf, d_u, d_z, flags=FL)
        e1.record()
        torch.cuda.synchronize()
        if it >= 2:
            ms_st += ex.stereo_ms()
            ms_all += e0.elapsed_time(e1)
    u = d_u[:nsc].cpu().numpy()
    counts = d_counts.cpu().numpy().reshape(P, 2)
    out = {"workload": "Frame::ComputeStereoMatches after extraction, %d rectified 752x480 pairs per step, %d features"
                       % (P, NFEAT), "stereo_kernels_ms_per_step": ms_st / steps, "call_ms_per_step": ms_all / steps,
           "pairs_per_s_stereo_kernels": P / (ms_st / steps * 1e-3) if ms_st > 0 else None,
           "pairs_per_s_call": P / (ms_all / steps * 1e-3),
           "matches_per_pair": float((u >= 0).sum() / nsc), "keypoints_left": float(counts[:, 0].mean())}
    # CPU oracle beside it (single thread, a few pairs), parity on those pairs
    sys.path.insert(0, os.path.join(ROOT, "oracle", "py"))
    import oracle_lib
    oracle_lib.build()
    t_cpu, ok = 0.0, True
    for k in range(min(nsc, 4)):
        eL, eR = (oracle_lib.OracleExtractor(NFEAT, 1.2, NLEVELS, 20, 7) for _ in range(2))
        _, kl, dl = eL(imgs[k, 0])
        _, kr, dr = eR(imgs[k, 1])
        t0 = time.perf_counter()
        ru, rz, _, _ = oracle_lib.stereo_matches(eL, eR, kl, dl, kr, dr, mb, mbf)
        t_cpu += time.perf_counter() - t0
        ok = ok and bool(np.array_equal(u[k, :len(kl)].view(np.uint32), ru.view(np.uint32)))
    out["cpu_oracle_pairs_per_s_1thread"] = min(nsc, 4) / t_cpu
    out["parity_sample"] = {"pairs": min(nsc, 4), "vs": "CPU oracle (bit-exact mvuRight)", "ok": ok}
    return out


def c4_leg(args, rank, world, local, dev, dist):
    """BASELINE configs[3]: 3840x2160 frames, 5000 features, 12 levels, a batch of frames per GPU (frames are
    independent: every rank extracts its own batch, no data-path collective; weak scaling).  Device-resident input and
    output, CUDA-event time of the call, max over ranks; a 4K frame is 26.8 Mpixel of pyramid (24x a C2 image)."""
    import torch
    from orbslam3lib_b200 import ORBextractor, _lib, scenes
    n, w, h, nf, nl = args.c4_frames, 3840, 2160, 5000, 12
    base = scenes.make("texture", w, h, 1004 + rank)
    imgs = np.empty((n, h, w), np.uint8)
    for k in range(n):
        imgs[k] = np.roll(base, 17 * k, axis=1)    # distinct frames without regenerating the scene
    d_imgs = torch.from_numpy(imgs).to(dev)
    ex = ORBextractor(nf, 1.2, nl, 20, 7, device=local)
    ex.set_stream(torch.cuda.current_stream().cuda_stream)
    cap = ex.max_keypoints
    d_kps = torch.zeros((n, cap, 28), dtype=torch.uint8, device=dev)
    d_desc = torch.zeros((n, cap, 32), dtype=torch.uint8, device=dev)
    d_counts = torch.zeros(n, dtype=torch.int32, device=dev)
    d_monos = torch.zeros(n, dtype=torch.int32, device=dev)
    FL = _lib.IN_DEVICE | _lib.OUT_DEVICE
    steps, ms, stages = 3, 0.0, {}
    for it in range(steps + 2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0.record()
        ex.extract_batch_raw(d_imgs, n, w, h, d_kps, d_desc, cap, d_counts, d_monos, flags=FL)
        e1.record()
        torch.cuda.synchronize()
        if it >= 2:
            ms += e0.elapsed_time(e1)
            for k, v in ex.stage_ms().items():
                stages[k] = stages.get(k, 0.0) + v
    ms /= steps
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    hbm_peak, _, _ = peaks()
    algo = {"pyramid": 45164255, "fast": 26804551, "blur": 53609102}    # SURVEY.md 8d, bytes per 4K frame
    rep = {}
    for k, v in stages.items():
        v /= steps
        rep[k] = {"ms_per_step": round(v, 4)}
        if k in algo and v > 0:
            rep[k]["frac_hbm"] = round(algo[k] * n / (v * 1e-3) * 1e-9 / hbm_peak, 4)
    return {"workload": "c4: %d frames 3840x2160 per GPU per step, 5000 features, 12 levels, %d GPU(s)" % (n, world),
            "ms_per_step": ms, "frames_per_s": n * world / (ms * 1e-3), "scaling": "weak",
            "keypoints_per_frame": float(d_counts.float().mean().item()), "stages": rep}


def search_leg(args, local, dev):
    """frame grid + ORBmatcher::SearchByProjection (ORBmatcher.cc:44-143) for a batch of frames
    (one CTA per frame), 2000 local-map points against ~1200 features per frame; everything resident in HBM.  The CPU
    oracle's orc_search_by_projection is timed beside it on one thread and checked on the sampled frames."""
    import torch
    from orbslam3lib_b200 import ORBextractor, ORBmatcher, _lib, scenes
    nfr, npts = args.search_frames, 2000
    ex = ORBextractor(NFEAT, 1.2, NLEVELS, 20, 7, device=local)
    cap = ex.max_keypoints
    nsc = min(nfr, 8)
    imgs = np.stack([scenes.make("texture", W, H, 7000 + k) for k in range(nsc)])
    counts, _, kps, desc = ex.extract_batch(imgs)
    scale = ex.GetScaleFactors()
    K = np.zeros((nfr, cap), _lib.KP_DTYPE)
    D = np.zeros((nfr, cap, 32), np.uint8)
    U = np.full((nfr, cap), -1, np.float32)
    P = np.zeros((nfr, npts), scenes.TRACK_DTYPE)
    nn = np.zeros(nfr, np.int32)
    for f in range(nfr):
        k = f % nsc
        n = int(counts[k])
        nn[f] = n
        K[f, :n], D[f, :n] = kps[k, :n], desc[k, :n]
        U[f, :n] = kps[k, :n]["x"] - 9.25
        if f < nsc:
            P[f] = scenes.track_points(kps[k, :n], desc[k, :n], U[f, :n], NLEVELS, npts, 7100 + f, dup_frac=0.2)
        else:
            P[f] = P[f % nsc]
    t = lambda a: torch.from_numpy(a.view(np.uint8).reshape(a.shape[0], -1)).to(dev)    # noqa: E731
    dK, dD, dU, dP = t(K), t(D), torch.from_numpy(U).to(dev), t(P)
    dn = torch.from_numpy(nn).to(dev)
    dnp = torch.full((nfr,), npts, dtype=torch.int32, device=dev)
    d_match = torch.zeros((nfr, npts), dtype=torch.int32, device=dev)
    d_owner = torch.zeros((nfr, cap), dtype=torch.int32, device=dev)
    d_cnt = torch.zeros(nfr, dtype=torch.int32, device=dev)
    m = ORBmatcher(0.8, device=local)
    bounds = (0.0, 0.0, float(W), float(H))
    FL = _lib.IN_DEVICE | _lib.OUT_DEVICE
    ms, steps = 0.0, 5
    for it in range(steps + 2):
        m.search_by_projection_batch_raw(dK, dD, dU, None, dn, cap, bounds, scale, dP, dnp, npts, nfr, 1.0, d_match,
                                         d_owner, d_cnt, flags=FL)
        if it >= 2:
            ms += m.last_ms()[0]
    match = d_match[:nsc].cpu().numpy()
    mpf = float(d_cnt.float().mean().item())
    sys.path.insert(0, os.path.join(ROOT, "oracle", "py"))
    import oracle_lib
    oracle_lib.build()
    ok, t_cpu = True, 0.0
    for f in range(nsc):
        n = int(nn[f])
        t0 = time.perf_counter()
        rm, _, _ = oracle_lib.search_by_projection(K[f, :n], D[f, :n], U[f, :n], None, bounds, scale, P[f], th=1.0,
                                                   nn_ratio=0.8)
        t_cpu += time.perf_counter() - t0
        ok = ok and bool(np.array_equal(match[f], rm))
    # ORBmatcher::SearchForInitialization (ORBmatcher.cc:649-764): frame k-1 -> frame k, window 100, ratio 0.9
    prev_imgs = np.stack([scenes.previous_frame(imgs[k], 7000 + k) for k in range(nsc)])
    pcounts, _, pkps, pdesc = ex.extract_batch(prev_imgs)
    K1 = np.zeros((nfr, cap), _lib.KP_DTYPE)
    D1 = np.zeros((nfr, cap, 32), np.uint8)
    PM = np.zeros((nfr, cap, 2), np.float32)
    n1 = np.zeros(nfr, np.int32)
    for f in range(nfr):
        k = f % nsc
        n = int(pcounts[k])
        n1[f] = n
        K1[f, :n], D1[f, :n] = pkps[k, :n], pdesc[k, :n]
        PM[f, :n, 0], PM[f, :n, 1] = pkps[k, :n]["x"], pkps[k, :n]["y"]
    dK1, dD1, dn1 = t(K1), t(D1), torch.from_numpy(n1).to(dev)
    dPM0 = torch.from_numpy(PM).to(dev)
    d_m12 = torch.zeros((nfr, cap), dtype=torch.int32, device=dev)
    mi = ORBmatcher(0.9, True, device=local)
    ms_i = 0.0
    for it in range(steps + 2):
        dPM = dPM0.clone()
        mi.search_for_initialization_batch_raw(dK1, dD1, dn1, dK, dD, dn, cap, nfr, bounds, 100, dPM, d_m12, d_cnt, flags=FL)
        if it >= 2:
            ms_i += mi.last_ms()[0]
    m12 = d_m12[:nsc].cpu().numpy()
    ok_i, t_cpu_i = True, 0.0
    for f in range(nsc):
        a, b = int(n1[f]), int(nn[f])
        t0 = time.perf_counter()
        rm, _, _ = oracle_lib.search_for_initialization(K1[f, :a], D1[f, :a], K[f, :b], D[f, :b], bounds, PM[f, :a], 100, 0.9,
                                                        True)
        t_cpu_i += time.perf_counter() - t0
        ok_i = ok_i and bool(np.array_equal(m12[f, :a], rm))
    init = {"workload": "ORBmatcher::SearchForInitialization, %d frame pairs per step, ~%d level-0 keypoints x window 100"
                        % (nfr, int((pkps[0, :int(pcounts[0])]["octave"] == 0).sum())),
            "kernel_ms_per_step": ms_i / steps, "pairs_per_s": nfr / (ms_i / steps * 1e-3),
            "matches_per_pair": float(d_cnt.float().mean().item()), "cpu_oracle_pairs_per_s_1thread": nsc / t_cpu_i,
            "parity_sample": {"pairs": nsc, "vs": "CPU oracle (vnMatches12)", "ok": ok_i}}
    # ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (ORBmatcher.cc:1683-1921): the per-frame tracking
    # call, 1200 last-frame points per frame, th = 7 (Tracking::TrackWithMotionModel, stereo), rotation histogram on
    nfp = 1200
    FP = np.zeros((nfr, nfp), scenes.FRAME_POINT_DTYPE)
    for f in range(nfr):
        k = f % nsc
        n = int(counts[k])
        FP[f] = scenes.frame_points(kps[k, :n], desc[k, :n], n, None, U[f, :n], NLEVELS, nfp, 7200 + f) if f < nsc else FP[f % nsc]
    dFP = t(FP)
    dnfp = torch.full((nfr,), nfp, dtype=torch.int32, device=dev)
    d_ml = torch.zeros((nfr, nfp), dtype=torch.int32, device=dev)
    d_mr = torch.zeros((nfr, nfp), dtype=torch.int32, device=dev)
    mf = ORBmatcher(0.9, True, device=local)
    ms_f = 0.0
    for it in range(steps + 2):
        mf.search_frame_batch_raw(dK, dD, dU, None, dn, None, None, cap, bounds, scale, 47.9, dFP, dnfp, nfp, nfr, 7.0, d_ml,
                                  d_mr, d_owner, d_cnt, flags=FL)
        if it >= 2:
            ms_f += mf.last_ms()[0]
    mlf = d_ml[:nsc].cpu().numpy()
    mpf_f = float(d_cnt.float().mean().item())
    ok_f, t_cpu_f = True, 0.0
    for f in range(nsc):
        n = int(nn[f])
        t0 = time.perf_counter()
        rl, _, _, _ = oracle_lib.search_by_projection_frame(K[f, :n], D[f, :n], U[f, :n], None, n, bounds, scale, 47.9, FP[f],
                                                            th=7.0)
        t_cpu_f += time.perf_counter() - t0
        ok_f = ok_f and bool(np.array_equal(mlf[f], rl))
    frame = {"workload": "ORBmatcher::SearchByProjection(CurrentFrame, LastFrame), %d frames per step, %d last-frame points x "
                         "~%d features, th = 7, rotation check" % (nfr, nfp, int(nn.mean())),
             "kernel_ms_per_step": ms_f / steps, "frames_per_s": nfr / (ms_f / steps * 1e-3), "matches_per_frame": mpf_f,
             "cpu_oracle_frames_per_s_1thread": nsc / t_cpu_f,
             "parity_sample": {"frames": nsc, "vs": "CPU oracle (slot per point after the rotation check)", "ok": ok_f}}
    return {"workload": "ORBmatcher::SearchByProjection, %d frames per step, %d map points x ~%d features per frame, th = 1"
                        % (nfr, npts, int(nn.mean())), "kernel_ms_per_step": ms / steps,
            "frames_per_s": nfr / (ms / steps * 1e-3), "matches_per_frame": mpf,
            "cpu_oracle_frames_per_s_1thread": nsc / t_cpu,
            "parity_sample": {"frames": nsc, "vs": "CPU oracle (match index per map point)", "ok": ok},
            "initialization": init, "frame_to_frame": frame}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle", "py"))
    import oracle_lib
    oracle_lib.build()
    cores = oracle_lib.hardware_threads()
    sample = max(4, min(32, cores // 2))
    imgs = make_pairs(sample, 1000)

    def step():
        counts, monos, kps, desc = oracle_lib.extract_batch(imgs, NFEAT, 1.2, NLEVELS, 20, 7, threads=cores)
        p = 0
        for k in range(sample):
            nq, nt = int(counts[k]), int(counts[sample + k])
            oracle_lib.knn2_hamming(desc[k, :nq], desc[sample + k, :nt], threads=cores)
            p += nq * nt
        return p
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    pairs = 0
    for _ in range(args.steps):
        pairs += step()
    dt = time.perf_counter() - t0
    fps = 2 * sample * args.steps / dt
    line = {"impl": "reference", "metric": "orb_frames_per_s", "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "c2: 752x480 stereo pairs, 1200 features, 8 levels, extraction + L->R Hamming kNN k=2",
                       "sample_pairs_per_step": sample},
            "knn_gpairs_per_s": pairs / dt * 1e-9,
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": "port",
                             "sample": "%d stereo pairs per step, all host threads (CPU oracle oracle/cpp = the reference's "
                                       "ORBextractor_old.cc restated with a multi-threaded batch entry; bit-identical to the "
                                       "reference's own build oracle/_ref, which has no batch entry)" % sample},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    line["cpu_baseline"].update(cpu_extra(cores))    # single-frame / 2-thread / matcher legs + the reference's own build
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--pairs", type=int, default=512, help="stereo pairs per GPU per step (512 = BASELINE configs[1])")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--map-rows", type=int, default=1000000, help="rows of the map-scale kNN leg (configs[4]); 0 = skip")
    ap.add_argument("--map-steps", type=int, default=10)
    ap.add_argument("--no-nf1000", dest="nf1000", action="store_false", help="skip the 1000-feature leg")
    ap.add_argument("--no-two-handles", dest="two_handles", action="store_false",
                    help="skip the e2e leg with two batches in flight (two extractor handles on two host threads)")
    ap.add_argument("--latency-iters", type=int, default=300, help="stereo pairs of the per-frame latency leg (configs[2]); 0 = skip")
    ap.add_argument("--c4-frames", type=int, default=32, help="4K frames per GPU per step of the configs[3] leg; 0 = skip")
    ap.add_argument("--search-frames", type=int, default=512, help="frames per step of the SearchByProjection leg; 0 = skip")
    ap.add_argument("--stereo-pairs", type=int, default=512, help="pairs per step of the ComputeStereoMatches leg; 0 = skip")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from orbslam3lib_b200 import BFMatcher, ORBextractor, _lib

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    P = args.pairs
    n = 2 * P
    imgs_np = make_pairs(P, 1000 + rank * P)    # every rank extracts its own frames: no data-path collective
    h_imgs = torch.from_numpy(imgs_np).pin_memory()
    d_imgs = h_imgs.to(dev)

    ex = ORBextractor(NFEAT, 1.2, NLEVELS, 20, 7, device=local)
    mt = BFMatcher(device=local)
    stream = torch.cuda.current_stream().cuda_stream
    ex.set_stream(stream)
    mt.set_stream(stream)
    cap = ex.max_keypoints
    u8 = torch.uint8
    d_kps = torch.zeros((n, cap, 28), dtype=u8, device=dev)
    d_desc = torch.zeros((n, cap, 32), dtype=u8, device=dev)
    d_counts = torch.zeros(n, dtype=torch.int32, device=dev)
    d_monos = torch.zeros(n, dtype=torch.int32, device=dev)
    d_idx = torch.zeros((P, cap, 2), dtype=torch.int32, device=dev)
    d_dist = torch.zeros((P, cap, 2), dtype=torch.int32, device=dev)
    h_kps = torch.zeros((n, cap, 28), dtype=u8).pin_memory()
    h_desc = torch.zeros((n, cap, 32), dtype=u8).pin_memory()
    h_counts = torch.zeros(n, dtype=torch.int32).pin_memory()
    h_monos = torch.zeros(n, dtype=torch.int32).pin_memory()
    h_idx = torch.zeros((P, cap, 2), dtype=torch.int32).pin_memory()
    h_dist = torch.zeros((P, cap, 2), dtype=torch.int32).pin_memory()
    FL = _lib.IN_DEVICE | _lib.OUT_DEVICE

    def step_device():
        ex.extract_batch_raw(d_imgs, n, W, H, d_kps, d_desc, cap, d_counts, d_monos, flags=FL)
        st = ex.stage_ms()
        mt.knn2_batch_raw(d_desc, d_counts, cap, d_desc[P:], d_counts[P:], cap, P, d_idx, d_dist, flags=FL)
        st["knn"] = mt.last_ms()[0]
        return st

    # e2e: ONE call of the public C ABI per step, host buffers in and out (orbfe_extract_match_batch = the batched form of
    # the Lynx wire call: both eyes + the left -> right matcher per pair; images interleaved L0, R0, L1, R1, ...)
    h_pairs = torch.from_numpy(np.ascontiguousarray(np.stack([imgs_np[:P], imgs_np[P:]], 1))).pin_memory()

    def step_e2e():
        ex.extract_match_batch_raw(h_pairs, P, W, H, h_kps, h_desc, cap, h_counts, h_monos, h_idx, h_dist)

    def step_e2e_two_calls():
        # the r1 form: extraction, then the matcher on the descriptors the extractor left in HBM
        ex.extract_batch_raw(h_imgs, n, W, H, h_kps, h_desc, cap, h_counts, h_monos, flags=0)
        _, dd, dc, dn, dcap = ex.device_outputs()
        assert dn == n and dcap == cap
        mt.knn2_batch_raw(dd, dc, cap, dd + P * cap * 32, dc + 4 * P, cap, P, h_idx, h_dist, flags=_lib.IN_DEVICE)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        acc = {}
        barrier()
        e0.record()
        for _ in range(steps):
            r = fn()
            if r:
                for k, v in r.items():
                    acc[k] = acc.get(k, 0.0) + v
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, acc

    for _ in range(args.warmup):
        step_device()
    sampler = ClockSampler(local)
    sampler.start()
    l0 = _lib.lib().orbfe_kernel_launches()
    ms_dev, stages = timed(step_device, args.steps)
    launches = int(_lib.lib().orbfe_kernel_launches() - l0)
    counts = d_counts.cpu().numpy().astype(np.int64)
    pairs_per_step = int((counts[:P] * counts[P:]).sum())
    for _ in range(max(1, args.warmup // 2)):
        step_e2e()
    ms_e2e, _ = timed(step_e2e, args.steps)
    # e2e with two batches in flight: a second extractor handle (own streams, own workspace, own pinned buffers) on a
    # second host thread; the two threads split the K steps, so the copy-in of one batch overlaps the kernel tail and
    # the copy-out of the other (handles are independent, include/orbfe_c.h; tests/test_gpu_threads.py).  Every step is
    # still one orbfe_extract_match_batch call with host buffers in and out.
    e2e_pair = None
    if args.two_handles:
        import threading
        ex_b = ORBextractor(NFEAT, 1.2, NLEVELS, 20, 7, device=local)
        s_b = torch.cuda.Stream(device=dev)
        ex_b.set_stream(s_b.cuda_stream)
        h_pairs_b = torch.flip(h_pairs, [0]).contiguous().pin_memory()    # the same pairs in reverse order
        b_kps, b_desc = torch.zeros_like(h_kps).pin_memory(), torch.zeros_like(h_desc).pin_memory()
        b_counts, b_monos = torch.zeros_like(h_counts).pin_memory(), torch.zeros_like(h_monos).pin_memory()
        b_idx, b_dist = torch.zeros_like(h_idx).pin_memory(), torch.zeros_like(h_dist).pin_memory()

        def step_b():
            ex_b.extract_match_batch_raw(h_pairs_b, P, W, H, b_kps, b_desc, cap, b_counts, b_monos, b_idx, b_dist)

        def run_pair(steps):
            errs = []

            def loop(fn, k):
                try:
                    for _ in range(k):
                        fn()
                except Exception as e:    # noqa: BLE001 - re-raised in the main thread
                    errs.append(e)
            ts = [threading.Thread(target=loop, args=(step_e2e, (steps + 1) // 2)),
                  threading.Thread(target=loop, args=(step_b, steps // 2))]
            for t in ts:
                t.start()
            for t in ts:
                t.join()
            if errs:
                raise errs[0]

        run_pair(2)
        ms_pair, _ = timed(lambda: run_pair(args.steps), 1)
        a_cnt = h_counts.numpy().reshape(P, 2)
        pair_same = bool(np.array_equal(b_counts.numpy().reshape(P, 2)[::-1], a_cnt) and
                         np.array_equal(b_idx.numpy()[::-1], h_idx.numpy()))
        e2e_pair = {"ms_per_step": ms_pair / args.steps, "equal_to_single_call": pair_same,
                    "value": n * world * args.steps / (ms_pair * 1e-3), "unit": "frames/s",
                    "call": "orbfe_extract_match_batch on two extractor handles from two host threads (two batches in flight)"}
        del ex_b
    clocks = sampler.stop()
    # after the timer: the fused call returns what the two-call form returns (pair p = images p and P + p there)
    e2e_idx, e2e_cnt = h_idx.numpy().copy(), h_counts.numpy().reshape(P, 2).copy()
    step_e2e_two_calls()
    torch.cuda.synchronize()
    ms_two, _ = timed(step_e2e_two_calls, max(2, args.steps // 2))
    e2e_same = bool(np.array_equal(e2e_idx, h_idx.numpy()) and np.array_equal(e2e_cnt[:, 0], h_counts.numpy()[:P]) and
                    np.array_equal(e2e_cnt[:, 1], h_counts.numpy()[P:]))

    # the configuration the metric string names (1000 features) beside c2 (1200): same images, device-resident
    nf1000 = None
    if args.nf1000:
        ex2 = ORBextractor(1000, 1.2, NLEVELS, 20, 7, device=local)
        ex2.set_stream(stream)
        cap2 = ex2.max_keypoints

        def step_1000():
            ex2.extract_batch_raw(d_imgs, n, W, H, d_kps, d_desc, cap2, d_counts, d_monos, flags=FL)
            dd = d_desc.view(-1)    # rows are cap2 apart in this leg
            mt.knn2_batch_raw(dd, d_counts, cap2, dd[P * cap2 * 32:], d_counts[P:], cap2, P, d_idx, d_dist, flags=FL)
        for _ in range(2):
            step_1000()
        ms_1000, _ = timed(step_1000, args.steps)
        nf1000 = {"workload": "752x480, 1000 features/image, 8 levels: extraction + L->R kNN, %d pairs per GPU per step" % P,
                  "ms_per_step": ms_1000 / args.steps, "frames_per_s": n * world * args.steps / (ms_1000 * 1e-3),
                  "keypoints_per_image": float(d_counts.float().mean().item())}
        del ex2
    knn_map = map_knn_leg(args, rank, world, local, dev, dist, mt) if args.map_rows > 0 else None
    c4 = None
    if args.c4_frames > 0:
        del d_imgs
        torch.cuda.empty_cache()
        c4 = c4_leg(args, rank, world, local, dev, dist)
    if world > 1:
        t = torch.tensor([float(pairs_per_step)], device=dev, dtype=torch.float64)
        dist.all_reduce(t)
        pairs_all = float(t.item())
    else:
        pairs_all = float(pairs_per_step)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    frames = n * world * args.steps
    fps = frames / (ms_dev * 1e-3)
    fps_e2e = frames / (ms_e2e * 1e-3)
    hbm_peak, peak_src, sm_mhz = peaks()
    per = {k: v / args.steps for k, v in stages.items()}    # ms per step per stage on rank 0
    total_stage = sum(per.values())
    stage_report = {}
    for k, v in per.items():
        e = {"ms_per_step": round(v, 4), "share": round(v / total_stage, 4) if total_stage else None}
        if k in ALGO_BYTES and v > 0:
            gbs = ALGO_BYTES[k] * n / (v * 1e-3) * 1e-9
            e["achieved_gbs"] = round(gbs, 1)
            e["frac_hbm"] = round(gbs / hbm_peak, 4)
        if k == "knn" and v > 0:
            gp = pairs_per_step / (v * 1e-3) * 1e-9
            e["gpairs_per_s"] = round(gp, 2)
            e["roofline"] = knn_roofline(gp, 1)
        stage_report[k] = e
    # pipe roofline: warp instructions per image and pipe from the committed ncu capture (profiles/kernel_traffic.json,
    # a property of kernel + input) x images / the CUDA-event time of THIS run, against the pipe's issue rate measured
    # in this process (orbfe_ubench_pipe)
    tp0 = os.path.join(ROOT, "profiles", "kernel_traffic.json")
    wi = json.load(open(tp0)).get("warp_inst_per_image", {}) if os.path.exists(tp0) else {}
    pipe_peak = {}
    for name, code in (("alu", 0), ("fmaheavy", 1)):
        import ctypes
        v = ctypes.c_double(0)
        if _lib.lib().orbfe_ubench_pipe(local, code, ctypes.byref(v)) == 0:
            pipe_peak[name] = v.value
    stage_kernels = {"pyramid": ["pad_interior_kernel", "pad_frame_kernel", "pad_level0_kernel", "resize_rows_kernel",
                                 "resize_tile_kernel"],
                     "fast": ["fast_cells_kernel"], "blur": ["gauss7x7_kernel"], "octree": ["octree_kernel"],
                     "describe": ["pack_kernel", "describe_kernel", "describe_fused_kernel", "describe_tma_kernel"]}
    for k, ks in stage_kernels.items():
        if k not in stage_report or per.get(k, 0) <= 0:
            continue
        best = None
        for pipe, peak in pipe_peak.items():
            cnt = sum(wi.get(kn, {}).get(pipe, 0.0) for kn in ks)
            if cnt <= 0:
                continue
            ach = cnt * n / (per[k] * 1e-3)
            e = {"pipe": pipe, "peak": peak, "achieved": ach, "frac": round(ach / peak, 4), "unit": "warp instructions/s",
                 "peak_source": "orbfe_ubench_pipe in this process", "count_source": "profiles/kernel_traffic.json warp_inst_per_image"}
            if best is None or e["frac"] > best["frac"]:
                best = e
        if best:
            tot = sum(wi.get(kn, {}).get("all", 0.0) for kn in ks)
            if tot > 0:
                best["issue_frac"] = round(tot * n / (per[k] * 1e-3) / (4 * 148 * sm_mhz * 1e6), 4)    # 4 schedulers x 148 SMs
            stage_report[k]["pipe_roofline"] = best
    hbm_stages = [k for k in ("pyramid", "fast", "blur") if per.get(k, 0) > 0]
    dom = max(hbm_stages, key=lambda k: per[k])
    launches_dom = NLEVELS if dom == "pyramid" else 1
    achieved = ALGO_BYTES[dom] * n / (per[dom] * 1e-3) * 1e-9
    traffic, traffic_src, issue_pct = None, None, None
    tp = os.path.join(ROOT, "profiles", "kernel_traffic.json")
    if os.path.exists(tp):    # dram bytes per image from the committed ncu --set full capture (not measured in this run)
        tall = json.load(open(tp))
        tj = tall["dram_bytes_per_image"]
        pad = ["pad_interior_kernel", "pad_frame_kernel"] if "pad_interior_kernel" in tj else ["pad_level0_kernel"]
        if dom == "pyramid":
            launches_dom = NLEVELS - 1 + len(pad)
        rs = ["resize_rows_kernel"] if "resize_rows_kernel" in tj else ["resize_tile_kernel"]
        ks = {"pyramid": pad + rs, "fast": ["fast_cells_kernel"], "blur": ["gauss7x7_kernel"]}[dom]
        if all(k in tj for k in ks):
            traffic = sum(tj[k] for k in ks) * n / launches_dom
            ip = tall.get("issue_active_pct", {})
            if all(k in ip for k in ks):
                issue_pct = max(ip[k] for k in ks)
            traffic_src = "profiles/kernel_traffic.json (ncu dram__bytes_read.sum + dram__bytes_write.sum per image x %d images / %d launches)" % (n, launches_dom)
    line = {
        "metric": "orb_frames_per_s", "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "c2: EuRoC-shaped 752x480 stereo pairs, 1200 features/image, 8 levels, scale 1.2, FAST 20/7, "
                               "%d pairs per GPU per step; extraction + left->right Hamming kNN k=2" % P,
                   "pairs_per_gpu": P, "l2_policy": "inputs larger than L2 (%.0f MB of images per step)" % (imgs_np.nbytes / 1e6),
                   "scene": "texture (seeded rectangles + box filter), right = row-dependent integer disparity"},
        "knn_gpairs_per_s": pairs_all * args.steps / (ms_dev * 1e-3) * 1e-9,
        "knn_pairs_per_step": pairs_all,
        "keypoints_per_image": float(counts.mean()),
        "e2e": {"value": fps_e2e, "unit": "frames/s", "ms_per_step": ms_e2e / args.steps,
                "call": "orbfe_extract_match_batch (one call: host images in; keypoints, descriptors, counts and k=2 matches out)",
                "two_call_ms_per_step": ms_two / max(2, args.steps // 2), "equals_two_call_form": e2e_same,
                "two_handles": e2e_pair,
                "h2d_bytes_per_step": int(h_imgs.numel()),
                "d2h_bytes_per_step": int(h_kps.numel() + h_desc.numel() + 2 * n * 4 + h_idx.numel() * 4 + h_dist.numel() * 4)},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                     "frac": achieved / hbm_peak, "traffic": traffic, "traffic_source": traffic_src,
                     "algorithmic_bytes_per_launch": ALGO_BYTES[dom] * n / launches_dom, "peak_source": peak_src,
                     "launches_per_step": launches_dom,
                     "issue_active_pct_ncu": issue_pct,    # committed ncu capture: why the kernel sits below the HBM roof
                     "note": "algorithmic bytes per image x images per step / CUDA-event time of the stage"},
        "stages": stage_report,
    }
    if nf1000:
        line["c2_nf1000"] = nf1000
    if knn_map:
        line["knn_map"] = knn_map
    if c4:
        line["c4_4k"] = c4
    if args.latency_iters > 0:
        line["latency"] = latency_leg(args, local)
    if args.stereo_pairs > 0:
        del d_kps, d_desc, d_idx, d_dist
        torch.cuda.empty_cache()
        line["stereo_matches"] = stereo_leg(args, local, dev)
    if args.search_frames > 0:
        torch.cuda.empty_cache()
        line["search_by_projection"] = search_leg(args, local, dev)
        torch.cuda.empty_cache()
        line["backend_matchers"] = backend_leg(args, local, dev)
    if not args.no_cpu_baseline:
        sys.path.insert(0, os.path.join(ROOT, "oracle", "py"))
        import oracle_lib
        cores = oracle_lib.hardware_threads()
        v, gp, sample = cpu_sample(max(4, min(32, cores // 2)), cores)
        line["cpu_baseline"] = {"value": v, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample,
                                "knn_gpairs_per_s": gp}
        line["cpu_baseline"].update(cpu_extra(cores))
    if knn_map:    # last key: survives a tail of the line
        line["knn_map_summary"] = {"n_gpus": world, "ms": round(knn_map["ms_per_step"], 3),
                                   "gpairs": round(knn_map["gpairs_per_s"], 1), "steps": knn_map["steps"],
                                   "speedup_vs_1gpu_ms": round(MAP_1GPU_MS / knn_map["ms_per_step"], 3),
                                   "ref_1gpu_ms": MAP_1GPU_MS, "comm_ms": knn_map.get("comm_ms"),
                                   "kernel_ms": knn_map.get("kernel_ms"), "checksum": knn_map["checksum"]}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
