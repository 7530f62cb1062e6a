#!/usr/bin/env python3
"""bench.py — frames/s of ORB extraction on KITTI-shaped 1241x376 frames, 2000 features (BASELINE.json metric).

A step = one pass of the extractor hot path over one batch of synthetic frames (default 512 per GPU).
  value  : whole-job frames/s with the batch already resident in HBM (orbx_extract_batch_device), timed with CUDA
           events on the launching stream, max over ranks.
  e2e    : the same metric through the host-buffer entry point orbx_extract_batch (pinned host frames in, keypoints +
           descriptors out, H2D and D2H inside the timed region).
  roofline: the dominant stage's algorithmic bytes / its CUDA-event duration, against MEASURED_PEAKS.json.
  cpu_baseline: the CPU oracle (port of the reference's CPU path) on the host cores, bounded sample, rank 0, N = 1.
`--impl reference` times the CPU oracle port alone (the reference's own sources do not compile here: they need
VisionWorks/OpenVX and OpenCV C++ headers).  Frames are sharded by frame across ranks with no
collective (weak scaling: the per-GPU batch is fixed).
"""
from __future__ import annotations

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

WORKLOADS = {  # reference YAMLs: Examples/Monocular/KITTI00-02.yaml, TUM1.yaml, Examples/Stereo/EuRoC.yaml
    "kitti": dict(shape=(376, 1241), nfeatures=2000, bytes_per_frame=7_770_771, sum_p=1_444_097),
    "tum": dict(shape=(480, 640), nfeatures=1000, bytes_per_frame=5_095_874, sum_p=950_532),
    "euroc": dict(shape=(480, 752), nfeatures=1200, bytes_per_frame=5_991_655, sum_p=1_117_367),
}
SCALE, LEVELS, INI_TH, MIN_TH = 1.2, 8, 20, 7
# roofline.traffic / roofline.issue come from profiles/ncu_latest.json, which tools/ncu_summary.py --json writes from an
# `ncu --set full` capture of one device-resident step (per workload; the file records the SHA-256 of the kernel sources it was
# taken from, and the line says whether they still match)
STAGE_KERNELS = {"Pyramid/Resize": ("k_pyramid_base", "k_pyramid_base_edges", "k_pyramid_resize4", "k_pyramid_resize"), "FAST+Grid": ("k_fast_strips",),
                 "Make quadtree": ("k_quadtree",), "Gaussian Blur": ("k_blur7",),
                 "Compute angle + ORB descriptor + Compute scale": ("k_orient_describe",)}


def ncu_latest(workload):
    import hashlib
    path = os.path.join(ROOT, "profiles", "ncu_latest.json")
    try:
        w = json.load(open(path))["workloads"][workload]
    except Exception:
        return None
    src = os.path.join(ROOT, "orb_slam_cuda_b200", "csrc")
    cur = {f: hashlib.sha256(open(os.path.join(src, f), "rb").read()).hexdigest()[:12] for f in w.get("kernel_sources_sha256_12", {}) if os.path.exists(os.path.join(src, f))}
    w["sources_match"] = cur == w.get("kernel_sources_sha256_12")
    return w


def ncu_stage(workload, stage, key):
    w = ncu_latest(workload)
    if not w:
        return None
    return sum(k[key] for n, k in w["kernels"].items() if n in STAGE_KERNELS.get(stage, ()))


METRIC = "frames/s ORB extract @1241x376 2k feats"


def issue_roofline(workload, frames_per_s_per_gpu, clocks):
    """The bound that actually holds for this integer path: warp instructions issued per second against 148 SMs x 4 schedulers x
    SM clock (one warp instruction per scheduler per cycle)."""
    w = ncu_latest(workload)
    if not w:
        return None
    per_frame = sum(k["warp_inst_per_frame"] for k in w["kernels"].values())
    sm_hz = ((clocks or {}).get("sm_mhz") or 1965.0) * 1e6
    peak = 148 * 4 * sm_hz
    ach = frames_per_s_per_gpu * per_frame
    return {"bound": "warp-instruction issue", "achieved": ach / 1e12, "peak": peak / 1e12, "unit": "T warp-inst/s", "frac": ach / peak,
            "warp_instructions_per_frame": per_frame,
            "source": f"ncu smsp__inst_executed.sum over the step's kernels (profiles/ncu_latest.json <- {w['report']}; sources match: {w['sources_match']})"}


def config_for(args):
    """The `config` object of both arms (the reference arm times a bounded sample of the same workload: cpu_baseline.sample)."""
    wl = WORKLOADS[args.workload]
    h, w = wl["shape"]
    return {"workload": f"{args.workload}: offline batch of {w}x{h} frames, {wl['nfeatures']} features, {LEVELS} levels x{SCALE}, FAST {INI_TH}/{MIN_TH}",
            "frames_per_gpu_per_step": args.batch, "sharding": "frame k -> rank k mod N, no collective",
            "l2": f"inputs {args.batch * w * h / 1e6:.0f} MB + intermediates per step exceed the 126 MB L2"}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="kitti", choices=list(WORKLOADS))
    ap.add_argument("--batch", type=int, default=512, help="frames per GPU per step")
    ap.add_argument("--cpu-frames", type=int, default=0, help="frames of the cpu_baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-threads", type=int, default=0, help="--impl reference: host threads (0 = all the process may use)")
    ap.add_argument("--no-match", action="store_true")
    ap.add_argument("--keyframes", type=int, default=1024, help="keyframes of the all-vs-all leg (config 5 names 4096: 16x the time)")
    ap.add_argument("--no-latency", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="profiling runs only: skip the host-buffer leg (the line then has no e2e)")
    ap.add_argument("--profile", action="store_true", help="profiling runs only: one warm-up and the timed steps, nothing else")
    return ap.parse_args()


def host_threads() -> int:
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx = float(r[2])
                for n, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def make_frames(workload: str, n: int, first: int = 0) -> np.ndarray:
    from orb_slam_cuda_b200 import synth
    return synth.frame_batch(workload, n, pool=min(n, 32), first=first)


def cpu_extract_rate(workload: str, nframes: int, threads: int, min_secs: float = 0.0):
    """frames/s of the CPU oracle, frame-parallel over `threads` host threads (one extractor per thread, as the
    reference runs one per eye, src/Frame.cc:77-80)."""
    from oracle import oracle_lib as O
    wl = WORKLOADS[workload]
    frames = make_frames(workload, nframes, first=1000)
    O.bench_extract(frames[:max(1, min(threads, nframes))], wl["nfeatures"], SCALE, LEVELS, INI_TH, MIN_TH, threads)  # warm
    secs, nk, done = 0.0, 0, 0
    while secs < min_secs or done == 0:  # the same block of frames again until the sample holds >= min_secs of CPU work
        t, k = O.bench_extract(frames, wl["nfeatures"], SCALE, LEVELS, INI_TH, MIN_TH, threads)
        secs, nk, done = secs + t, nk + k, done + nframes
    return done / secs, secs, done


def match_leg(X, torch, dist, rank, world, stream, K, args_keyframes=1024):
    """Gmatch/s (descriptor pairs compared per second / 1e9) of the Hamming matchers, device resident, CUDA events on
    the launching stream: (a) 2000 x 2000 brute force best / second-best + TH_LOW / ratio test (config 4's inner loop),
    (b) all-vs-all keyframe matching sharded by query block, database all-gathered over NCCL when world > 1
    (config 5).  Strong scaling for (b): the keyframe count is fixed as N grows."""
    from orb_slam_cuda_b200 import synth, sharding
    m = X.ORBmatcher(0.6, True)
    q, d = synth.descriptor_sets(4, 2000, 2000)
    dq, dd = torch.from_numpy(q).cuda(), torch.from_numpy(d).cuda()
    out = [torch.empty(2000, dtype=torch.int32, device="cuda") for _ in range(4)]
    reps = 200

    def bf():
        m.match_bruteforce_device(dq.data_ptr(), 2000, dd.data_ptr(), 2000, *[o.data_ptr() for o in out], stream=stream)

    for _ in range(10):
        bf()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        bf()
    e1.record()
    torch.cuda.synchronize()
    bf_ms = e0.elapsed_time(e1) / reps
    # (b) all-vs-all (config 5): k_total keyframes x 2000 descriptors, query block per rank, database all-gathered over NCCL
    # inside liborbx.so (orbx_allpairs_sharded: own block first, chunked ncclAllGather overlapped with the scoring).
    k_total, n = args_keyframes, 2000
    b, e = sharding.block_range(k_total, rank, world)
    rng = np.random.default_rng(77)
    pool = rng.integers(0, 256, size=(64, n, 32), dtype=np.uint8)  # 64 distinct keyframes, every further one a row rotation + bit noise
    def keyframe(i):
        kf = np.roll(pool[i % 64], i // 64, axis=0)
        if i >= 64:
            kf = kf ^ (np.random.default_rng(i).random((n, 32)) < 0.02).astype(np.uint8)
        return kf
    local = torch.from_numpy(np.stack([keyframe(i) for i in range(b, e)]) if e > b else np.zeros((0, n, 32), np.uint8)).cuda()
    comm = sharding.NcclComm(torch.device("cuda", torch.cuda.current_device()))
    scores = torch.empty((e - b, k_total), dtype=torch.int32, device="cuda")

    def ap():
        return sharding.allpairs_sharded_native(comm, local, k_total, 50, 0.6, chunks=4, stream=stream, scores=scores)

    ap()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(K):
        sc, moved = ap()
    e1.record()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / K], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ap_ms = float(t[0])
    # ---- outside the timed region: the sharded result is checked on every rank ----
    # (1) a sample of query rows of this rank against the single-GPU kernel over the FULL database, built locally (every rank
    #     can regenerate every keyframe), (2) on rank 0 a sample of entries against the CPU oracle's brute-force matcher
    check = {"rows_vs_single_gpu": 0, "entries_vs_cpu_oracle": 0, "ok": True}
    if e > b:
        full = torch.from_numpy(np.stack([keyframe(i) for i in range(k_total)])).cuda()
        rows = sorted(set([0, (e - b) // 2, e - b - 1]))
        ref = torch.empty((len(rows), k_total), dtype=torch.int32, device="cuda")
        m.match_allpairs_device(local[rows].contiguous().data_ptr(), len(rows), full.data_ptr(), k_total, n, ref.data_ptr(), stream=stream)
        torch.cuda.synchronize()
        check["rows_vs_single_gpu"] = len(rows)
        check["ok"] = bool(torch.equal(ref, sc[rows]))
        if rank == 0 and check["ok"]:
            from oracle import oracle_lib as O  # the checker, outside the timed region
            sc_h = sc.cpu().numpy()
            for (qi, dj) in [(0, b), (0, (b + 64) % k_total), ((e - b) // 2, (b + (e - b) // 2 + 128) % k_total), (e - b - 1, 3 % k_total)]:
                _, d1, d2, mt = O.match_bruteforce(keyframe(b + qi), keyframe(dj), 50, 0.6)
                check["ok"] = check["ok"] and int((mt >= 0).sum()) == int(sc_h[qi, dj])
                check["entries_vs_cpu_oracle"] += 1
        del full
    ok = torch.tensor([1 if check["ok"] else 0], dtype=torch.int32, device="cuda")
    if dist is not None:
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    check["ok"] = bool(int(ok[0]) == 1)
    comm.destroy()
    if not check["ok"]:
        raise SystemExit("bench.py: the sharded all-vs-all scores differ from the single-GPU kernel / the CPU oracle")
    return {"bruteforce_2000x2000": {"gmatch_per_s": 2000 * 2000 / (bf_ms * 1e-3) / 1e9, "ms": bf_ms, "launches_per_call": 1},
            "allpairs": {"gmatch_per_s": float(k_total) * k_total * n * n / (ap_ms * 1e-3) / 1e9, "ms": ap_ms,
                         "keyframes": k_total, "descriptors_per_keyframe": n,
                         "sharding": "query block per rank; orbx_allpairs_sharded: own block first, database ncclAllGather'ed in 4 chunks on a second stream",
                         "allgather_bytes_per_rank": int(moved), "scaling": "strong", "steps": K, "check": check},
            "unit": "Gmatch/s (descriptor pairs compared / s / 1e9)"}


def latency_leg(X, wl, w, h, kind, cpu=False):
    """p50 / p95 wall-clock latency of single-frame orbx_extract calls (pageable host frame in, keypoints + descriptors
    out; what ORBextractor::operator() costs the Tracking thread per frame)."""
    from orb_slam_cuda_b200 import synth
    ex1 = X.ORBextractor(wl["nfeatures"], SCALE, LEVELS, INI_TH, MIN_TH, w, h, max_batch=1)
    imgs = [synth.frame(kind, 500 + i) for i in range(8)]
    for i in range(10):
        ex1(imgs[i % 8])
    ts = []
    for i in range(200):
        t0 = time.perf_counter()
        ex1(imgs[i % 8])
        ts.append((time.perf_counter() - t0) * 1e3)
    ts = np.sort(np.array(ts))
    out = {"p50_ms": float(ts[len(ts) // 2]), "p95_ms": float(ts[int(len(ts) * 0.95)]), "calls": len(ts),
           "path": "orbx_extract through the python binding, pageable host buffers"}
    # the same call as a C++ caller of the class surface makes it (ORBextractor::operator(), tests/cpp/frame_harness latency):
    # what a maintainer of the reference sees per frame, without the Python binding's share
    harness = os.path.join(ROOT, "tests", "cpp", "frame_harness")
    if os.path.exists(harness) and not os.environ.get("CUDA_INJECTION64_PATH"):  # not under ncu / another injected profiler
        import subprocess
        import tempfile
        try:
            with tempfile.NamedTemporaryFile(suffix=".raw") as f:
                f.write(imgs[0].tobytes() + imgs[1].tobytes())
                f.flush()
                env = dict(os.environ, LD_LIBRARY_PATH=os.path.join(ROOT, "orb_slam_cuda_b200") + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
                r = subprocess.run([harness, "latency", f.name, str(w), str(h), str(wl["nfeatures"]), "300"], capture_output=True, text=True,
                                   timeout=120, env=env)
            j = json.loads(r.stdout.strip().splitlines()[-1])
            out["cpp_class_surface"] = {"p50_ms": j["p50_us"] / 1e3, "p95_ms": j["p95_us"] / 1e3, "calls": j["calls"],
                                        "path": "ORBextractor::operator() of include/ORBextractor.h from C++ (tests/cpp/frame_harness latency)"}
        except Exception as e:  # the Python figure above stands on its own
            out["cpp_class_surface"] = {"unavailable": repr(e)[:200]}

    def p50(fn, n=100, warm=5):
        for _ in range(warm):
            fn()
        t = []
        for _ in range(n):
            t0 = time.perf_counter()
            fn()
            t.append((time.perf_counter() - t0) * 1e3)
        return float(np.sort(np.array(t))[n // 2])

    # config 3: EuRoC-shaped stereo pair, 1200 features per eye, one extractor per eye on two host threads
    # (src/Frame.cc:77-80), then ComputeStereoMatches on the device-resident pyramids
    left, right = synth.stereo_pair("euroc", 9)
    exL, exR = (X.ORBextractor(1200, SCALE, LEVELS, INI_TH, MIN_TH, 752, 480, max_batch=1) for _ in range(2))
    res = {}

    def eye(name, e, img):
        res[name] = e(img)

    from concurrent.futures import ThreadPoolExecutor
    pool = ThreadPoolExecutor(2)  # the two eye threads live across frames (a thread start per frame costs ~50 us in Python)

    def stereo_frame():
        fl, fr = pool.submit(eye, "L", exL, left), pool.submit(eye, "R", exR, right)
        fl.result(); fr.result()
        (kl, dl), (kr, dr) = res["L"], res["R"]
        res["u"] = X.ComputeStereoMatchesLast(exL, exR, len(kl), 47.90639384423901, 435.2046959714599)  # on the device-resident output

    out["stereo_euroc"] = {"p50_ms": p50(stereo_frame), "what": "left + right extraction on two (pooled) host threads + ComputeStereoMatches on the extractors' device-resident output (orbx_stereo_match_last), 752x480, 1200 features per eye"}
    out["stereo_euroc"]["matched"] = int((res["u"][0] >= 0).sum())
    # config 4: frame-to-frame matching of two 2000-feature KITTI frames: SearchForInitialization (window 100, ratio 0.9)
    a, b = synth.frame("kitti", 40), synth.shifted_frame("kitti", 40)
    ex2 = X.ORBextractor(2000, SCALE, LEVELS, INI_TH, MIN_TH, 1241, 376, max_batch=1)
    (k1, d1), (k2, d2) = ex2(a), ex2(b)
    k1, d1, k2, d2 = k1.copy(), d1.copy(), k2.copy(), d2.copy()
    m = X.ORBmatcher(0.9, True)
    bounds = (0.0, 1241.0, 0.0, 376.0)
    sfi = {}

    def run_sfi():
        prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
        sfi["n"], _ = m.SearchForInitialization(k1, d1, k2, d2, bounds, prev, 100)

    out["search_for_initialization_2000x2000"] = {"p50_ms": p50(run_sfi), "matches": int(sfi["n"]),
                                                   "what": "host buffers in and out, window 100, TH_LOW 50, ratio 0.9, rotation check"}
    out["tracking"] = tracking_leg(X, k1, d1, k2, d2, bounds, p50, cpu)
    return out


def orbvoc_shaped_vocabulary(k=10, L=6, seed=3):
    """A synthetic tree with ORBvoc.txt's shape (k = 10, L = 6: 1 111 110 nodes, 10^6 words, 35 MB of descriptors), numbered level
    by level; a child's descriptor is its parent's with 96 / level random bits flipped."""
    rng = np.random.default_rng(seed)
    parent, desc = [np.array([-1], np.int32)], [np.zeros((1, 32), np.uint8)]
    base, n = 0, 1
    for level in range(1, L + 1):
        par = base + np.repeat(np.arange(n, dtype=np.int32), k)
        if level == 1:
            d = rng.integers(0, 256, (n * k, 32), dtype=np.uint8)
        else:
            nb = max(96 // level, 8)
            flips = np.zeros((n * k, 32), np.uint8)
            pos = rng.integers(0, 256, (n * k, nb))
            np.bitwise_or.at(flips, (np.repeat(np.arange(n * k), nb), (pos >> 3).ravel()), (1 << (pos & 7)).astype(np.uint8).ravel())
            d = np.repeat(desc[-1], k, axis=0) ^ flips
        parent.append(par)
        desc.append(d)
        base += n
        n *= k
    parent, desc = np.concatenate(parent), np.concatenate(desc)
    leaf = np.zeros(parent.size, np.uint8)
    leaf[-n:] = 1
    weight = np.zeros(parent.size, np.float64)
    weight[-n:] = rng.uniform(0.5, 12.0, n)
    return k, L, parent, leaf, desc, weight


def tracking_leg(X, k1, d1, k2, d2, bounds, p50, cpu):
    """The per-frame matchers that follow extraction in Tracking on two 2000-feature KITTI frames:
    SearchByProjection (TrackWithMotionModel, src/Tracking.cc:962), Frame::ComputeBoW and SearchByBoW (TrackReferenceKeyFrame, :842).
    cpu_port_ms = the oracle's restatement of the same routine on one host thread (what the Tracking thread pays in the reference)."""
    out = {}
    n1, n2 = k1.size, k2.size
    cam = (718.856, 718.856, 607.1928, 185.2157)
    z = np.full(n1, 10.0, np.float32)
    x3 = np.stack([(k1["x"] + 7.0 - cam[2]) * z / cam[0], (k1["y"] - 4.0 - cam[3]) * z / cam[1], z], 1).astype(np.float32)
    one1, zero2 = np.ones(n1, np.uint8), np.zeros(n2, np.uint8)
    sf = (np.float32(SCALE) ** np.arange(LEVELS, dtype=np.float32)).astype(np.float32)
    m = X.ORBmatcher(0.9, True)
    r = {}

    def run_sbp():
        r["sbp"] = m.SearchByProjection(x3, one1, one1, d1, k1["octave"], k1["angle"], k2, d2, None, zero2, cam, 386.1448, bounds, sf,
                                        15.0, 0)

    out["search_by_projection_2000"] = {"p50_ms": p50(run_sbp), "matches": int(r["sbp"][0]),
                                        "what": "frame-to-frame, th 15, mono, TH_HIGH 100, rotation check; host buffers in and out"}
    # the same call with the current frame resident on the device (orbx_frame: keypoints, descriptors and the 64 x 48 grid uploaded and
    # binned once per frame, as Tracking matches against the current frame several times)
    fr = X.DeviceFrame(k2, d2, bounds)

    def run_sbp_h():
        r["sbph"] = m.SearchByProjectionFrame(x3, one1, one1, d1, k1["octave"], k1["angle"], fr, zero2, cam, 386.1448, sf, 15.0, 0)

    out["search_by_projection_2000"]["p50_ms_frame_handle"] = p50(run_sbp_h)
    assert r["sbph"][0] == r["sbp"][0] and np.array_equal(r["sbph"][1], r["sbp"][1])
    k, L, parent, leaf, desc, weight = orbvoc_shaped_vocabulary()
    voc = X.ORBVocabulary(k, L, 0, 0, parent, leaf, desc, weight)
    # features that descend like real ones: leaf descriptors of the tree with a few bits flipped
    rng = np.random.default_rng(8)
    f1 = desc[-rng.integers(1, 10 ** L, n1)] ^ (rng.integers(0, 256, (n1, 32), dtype=np.uint8) & rng.integers(0, 256, (n1, 32), dtype=np.uint8)
                                                 & rng.integers(0, 256, (n1, 32), dtype=np.uint8) & rng.integers(0, 256, (n1, 32), dtype=np.uint8))
    f2 = f1[rng.permutation(n1)[:n2] if n2 <= n1 else rng.integers(0, n1, n2)] ^ (1 << rng.integers(0, 8, (n2, 32))).astype(np.uint8) * (rng.random((n2, 32)) < 0.1)

    def run_bow():
        r["bow"] = voc.transform_features(f1, 4)

    out["compute_bow_2000"] = {"p50_ms": p50(run_bow), "words": int(np.unique(r["bow"][0]).size),
                               "what": "DBoW2 descent of 2000 descriptors, synthetic tree of ORBvoc's shape (k 10, L 6, 1.1 M nodes), levelsup 4; "
                                       "device part of Frame::ComputeBoW (the std::map fills stay on the host)"}

    def node_lists(node):
        order = np.argsort(node, kind="stable")
        ids, start = np.unique(node[order], return_index=True)
        return ids, np.append(start, node.size).astype(np.int32), order.astype(np.uint32)

    na, sa, ia = node_lists(voc.transform_features(f1, 4)[2])
    nb, sb, ib = node_lists(voc.transform_features(f2, 4)[2])
    common, pa, pb = np.intersect1d(na, nb, return_indices=True)
    s1 = np.concatenate([[0], np.cumsum(sa[pa + 1] - sa[pa])]).astype(np.int32)
    s2 = np.concatenate([[0], np.cumsum(sb[pb + 1] - sb[pb])]).astype(np.int32)
    i1 = np.concatenate([ia[sa[p]:sa[p + 1]] for p in pa]) if common.size else np.zeros(0, np.uint32)
    i2 = np.concatenate([ib[sb[p]:sb[p + 1]] for p in pb]) if common.size else np.zeros(0, np.uint32)
    a1 = np.resize(k1["angle"], n1).astype(np.float32)
    a2 = np.resize(k2["angle"], n2).astype(np.float32)
    mb = X.ORBmatcher(0.7, True)
    one2 = np.ones(n2, np.uint8)

    def run_sbb():
        r["sbb"] = mb.SearchByBoW(s1, i1, s2, i2, one1, f1, a1, one2, f2, a2, 50)

    out["search_by_bow_2000"] = {"p50_ms": p50(run_sbb), "matches": int(r["sbb"][0]), "common_nodes": int(common.size),
                                 "what": "keyframe vs frame over the common level-2 nodes, TH_LOW 50, ratio 0.7, rotation check; host buffers"}
    if cpu:
        from oracle import oracle_lib as O  # the CPU port, timed as the baseline of these rows

        def t_ms(fn, n=5):
            fn()
            t = []
            for _ in range(n):
                t0 = time.perf_counter()
                fn()
                t.append((time.perf_counter() - t0) * 1e3)
            return float(np.sort(np.array(t))[n // 2])

        out["search_by_projection_2000"]["cpu_port_ms"] = t_ms(lambda: O.search_by_projection_frame(
            x3, one1, one1, d1, k1["octave"], k1["angle"], k2, d2, None, zero2, cam, 386.1448, bounds, sf, 15.0, 0, True))
        ov = O.Vocabulary(k, L, 0, 0, parent, leaf, desc, weight)
        # the oracle rebuilds its child lists on every call; a 1-feature call measures that and is subtracted
        out["compute_bow_2000"]["cpu_port_ms"] = max(t_ms(lambda: ov.transform_features(f1, 4), 5) - t_ms(lambda: ov.transform_features(f1[:1], 4), 5), 0.0)
        out["search_by_bow_2000"]["cpu_port_ms"] = t_ms(lambda: O.search_by_bow(s1, i1, s2, i2, one1, f1, a1, one2, f2, a2, 50, 0.7, True))
    return out


def run_reference(args, emit=print):
    """--impl reference: the reference's CPU ORBextractor path on all host threads.  kind "reference" = the reference's own
    src/ORBextractor.cc object code (oracle/_ref/libref_extractor.so, built by __graft_entry__.build() where /root/reference
    exists and carried to the GPU box; its OpenCV pixel primitives are scalar C++ models, OpenCV itself is not on the image);
    kind "port" = the oracle's restatement, when that library is absent."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = WORKLOADS[args.workload]
    threads = args.cpu_threads or host_threads()
    per_step = max(64, 16 * threads) if threads > 1 else 24  # bounded sample of the workload per step
    from oracle.ref_extractor import ref_extractor as RX
    if os.path.exists(RX.SO) and not os.environ.get("ORBX_BENCH_CPU_PORT"):
        kind, bench_extract = "reference", RX.bench_extract
        what = "reference src/ORBextractor.cc object code (CPU branch; OpenCV primitives = scalar C++ models)"
    else:
        from oracle import oracle_lib as O
        kind, bench_extract = "port", O.bench_extract
        what = "oracle restatement"
    frames = make_frames(args.workload, per_step, first=1000)
    a = (wl["nfeatures"], SCALE, LEVELS, INI_TH, MIN_TH, threads)
    for _ in range(max(args.warmup, 1)):
        bench_extract(frames[:threads], *a)
    t = 0.0
    for _ in range(args.steps):
        secs, _ = bench_extract(frames, *a)
        t += secs
    value = per_step * args.steps / t
    h, w = wl["shape"]
    line = {
        "impl": "reference", "metric": METRIC if args.workload == "kitti" else f"frames/s ORB extract ({args.workload})",
        "value": value, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": config_for(args),
        "cpu_baseline": {"value": value, "unit": "frames/s", "cores": threads, "kind": kind,
                         "sample": f"{per_step} frames per step x {args.steps} steps, frame-parallel over {threads} threads; {what}"},
        "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(json.dumps(line))


def run_ours(args, emit=print):
    import torch
    import orb_slam_cuda_b200 as X

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    wl = WORKLOADS[args.workload]
    h, w = wl["shape"]
    B, K, W = args.batch, args.steps, (max(args.warmup, 1) if args.profile else max(args.warmup, 3))
    cap = wl["nfeatures"] + 4 * LEVELS
    ex = X.ORBextractor(wl["nfeatures"], SCALE, LEVELS, INI_TH, MIN_TH, w, h, max_batch=B, device=local)

    # frames of this rank's shard (frame k of the job -> rank k mod world): distinct images, > L2 in total
    frames_np = make_frames(args.workload, B, first=rank * B)
    h_frames = torch.from_numpy(frames_np).pin_memory()
    d_frames = h_frames.cuda(non_blocking=True)
    d_kps = torch.empty((B, cap, 7), dtype=torch.int32, device="cuda")
    d_desc = torch.empty((B, cap, 32), dtype=torch.uint8, device="cuda")
    d_counts = torch.empty(B, dtype=torch.int32, device="cuda")
    d_status = torch.zeros(1, dtype=torch.int32, device="cuda")
    tstream = torch.cuda.Stream()  # a real (non-default) stream: CUDA events below are recorded on it
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream
    assert stream != 0

    def step_device():
        ex.extract_batch_device(d_frames.data_ptr(), B, w, h, d_kps.data_ptr(), d_desc.data_ptr(), cap, d_counts.data_ptr(),
                                d_status.data_ptr(), stream)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput ----
    for _ in range(W):
        step_device()
    barrier()
    launches0 = X.kernel_launches()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        step_device()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    launches = X.kernel_launches() - launches0
    clocks = sampler.stop() if rank == 0 else None
    status = int(d_status.item())
    counts = d_counts.cpu().numpy()

    if args.profile:  # under ncu: the steps above are all that runs
        if rank == 0:
            emit(json.dumps({"profile_run": True, "ms_per_step": ms / K, "status": status}))
        return
    # ---- per-stage times (CUDA events between the stages, same stream), averaged over K more steps ----
    ex.enable_timing(True)
    stage = {}
    for _ in range(K):
        step_device()
        for name, v in ex.stage_ms().items():
            stage[name] = stage.get(name, 0.0) + v / K
    ex.enable_timing(False)

    # ---- end to end through the host-buffer entry points: every step copies its frames from pinned host memory to the device
    #      and its keypoints + descriptors + counts back, inside the timed region.  Steps are enqueued with
    #      orbx_extract_batch_async and collected with orbx_extract_batch_wait, two in flight (two sets of output buffers), so
    #      step k + 1's H2D overlaps step k's kernels and D2H — what an offline sequence (config 5) does.  e2e_sync is the
    #      same through the blocking orbx_extract_batch (one step at a time). ----
    outs = []
    for _ in range(2):
        hk = torch.empty((B, cap, 7), dtype=torch.int32).pin_memory()
        hd = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
        hc = torch.empty(B, dtype=torch.int32).pin_memory()
        outs.append((hk.numpy().view(X.KP_DTYPE).reshape(B, cap), hd.numpy(), hc.numpy()))
    frames_host = h_frames.numpy()

    def step_host():
        ex.extract_batch(frames_host, *outs[0], cap)

    def run_pipelined(n):
        ex.extract_batch_async(frames_host, *outs[0], cap)
        for k in range(1, n):
            ex.extract_batch_async(frames_host, *outs[k & 1], cap)
            ex.extract_batch_wait()
        ex.extract_batch_wait()

    e2e_s = e2e_sync_s = float("nan")
    h2d_gbs = d2h_gbs = duplex_ms = float("nan")
    if not args.no_e2e:
        # what the box gives this rank while every rank copies at once: the bare pinned-host -> device and device -> pinned-host
        # rates of the step's own buffers (no kernels).  e2e cannot beat frames_per_step / (H2D time) on this box.
        d_tmp = torch.empty((B, cap * 60 // 4), dtype=torch.int32, device="cuda")  # keypoints + descriptors of a step
        h_tmp = torch.empty((B, cap * 60 // 4), dtype=torch.int32).pin_memory()
        s_up, s_down = torch.cuda.Stream(), torch.cuda.Stream()
        rates = {}
        for direction in ("h2d", "d2h", "duplex"):
            barrier()
            c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            c0.record()
            s_up.wait_stream(torch.cuda.current_stream())
            s_down.wait_stream(torch.cuda.current_stream())
            for _ in range(5):
                if direction in ("h2d", "duplex"):
                    with torch.cuda.stream(s_up):
                        d_frames.copy_(h_frames, non_blocking=True)
                if direction in ("d2h", "duplex"):
                    with torch.cuda.stream(s_down):
                        h_tmp.copy_(d_tmp, non_blocking=True)
            torch.cuda.current_stream().wait_stream(s_up)
            torch.cuda.current_stream().wait_stream(s_down)
            c1.record()
            barrier()
            rates[direction] = c0.elapsed_time(c1) / 5  # ms per step's worth of copies
        h2d_gbs = B * w * h / (rates["h2d"] * 1e-3) / 1e9
        d2h_gbs = d_tmp.numel() * 4 / (rates["d2h"] * 1e-3) / 1e9
        duplex_ms = rates["duplex"]
        del d_tmp, h_tmp
        for _ in range(W):
            step_host()
        barrier()
        t0 = time.perf_counter()
        for _ in range(K):
            step_host()
        barrier()
        e2e_sync_s = time.perf_counter() - t0
        assert np.array_equal(outs[0][2], counts), "host and device paths disagree"
        run_pipelined(W)
        barrier()
        t0 = time.perf_counter()
        run_pipelined(K)
        barrier()
        e2e_s = time.perf_counter() - t0
        d_k, d_d = d_kps.cpu().numpy(), d_desc.cpu().numpy()
        for hk_np, hd_np, hc_np in outs[:min(K, 2)]:  # both buffer sets hold the device path's result, bit for bit
            assert np.array_equal(hc_np, counts), "pipelined host path and device path disagree (counts)"
            for f in (0, B // 2, B - 1):
                n = int(counts[f])
                assert np.array_equal(np.ascontiguousarray(hk_np[f, :n]).view(np.int32).reshape(n, 7), d_k[f, :n]), "keypoints differ"
                assert np.array_equal(hd_np[f, :n], d_d[f, :n]), "descriptors differ"

    match = None if args.no_match else match_leg(X, torch, dist, rank, world, stream, 2, args.keyframes)

    t = torch.tensor([ms, e2e_s * 1e3, e2e_sync_s * 1e3, -h2d_gbs, -d2h_gbs, duplex_ms], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms, e2e_sync_ms, h2d_gbs, d2h_gbs, duplex_ms = (float(t[0]), float(t[1]), float(t[2]), -float(t[3]), -float(t[4]),
                                                            float(t[5]))  # copy rates: slowest rank
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    total_frames = world * B * K
    value = total_frames / (ms * 1e-3)
    peak, peak_src = peaks()
    dom = max(stage, key=stage.get)
    stage_bytes = {  # algorithmic bytes per frame of each stage
        "Pyramid/Resize": wl["bytes_per_frame"] - 4 * wl["sum_p"] - 60 * wl["nfeatures"] + wl["sum_p"],
        "FAST+Grid": wl["sum_p"],
        "Gaussian Blur": 2 * wl["sum_p"],
        "Make quadtree": 0,
        "Compute angle + ORB descriptor + Compute scale": 60 * wl["nfeatures"],
    }
    dom_bytes = stage_bytes.get(dom, 0) * B
    achieved = dom_bytes / (stage[dom] * 1e-3) / 1e9 if stage[dom] > 0 else 0.0
    line = {
        "metric": METRIC if args.workload == "kitti" else f"frames/s ORB extract ({args.workload})",
        "value": value, "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": config_for(args), "mean_keypoints": float(counts.mean()),
        "clocks": clocks,
        "e2e": {"value": total_frames / (e2e_ms * 1e-3), "unit": "frames/s", "h2d_bytes_per_step": B * w * h,
                "d2h_bytes_per_step": B * cap * 60 + B * 4 + 4, "ms_per_step": e2e_ms / K,
                "path": "orbx_extract_batch_async / orbx_extract_batch_wait, two steps in flight, pinned host buffers",
                "h2d_ceiling_gbs": h2d_gbs, "d2h_ceiling_gbs": d2h_gbs,
                "h2d_bound_frames_per_s": world * h2d_gbs * 1e9 / (w * h),
                "copies_only_ms_per_step": duplex_ms, "copy_bound_frames_per_s": world * B / (duplex_ms * 1e-3),
                "ceiling": "bare copies of the step's own pinned buffers (h2d: the frames; d2h: keypoints + descriptors; copies_only: both "
                           "directions at once on two streams), all ranks at once, slowest rank, no kernels: what the box allows",
                "sync": {"value": total_frames / (e2e_sync_ms * 1e-3), "ms_per_step": e2e_sync_ms / K,
                         "path": "blocking orbx_extract_batch, one step at a time"}},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": (ncu_stage(args.workload, dom, "dram_bytes_per_frame") or 0) * B or None,
                     "traffic_source": "ncu dram__bytes_read + dram__bytes_write of this stage's kernels per frame x frames per launch "
                                       f"(profiles/ncu_latest.json; sources match: {(ncu_latest(args.workload) or {}).get('sources_match')})",
                     "peak_source": peak_src, "algorithmic_bytes_per_launch": dom_bytes,
                     "avg_launch_ms": stage[dom],
                     "pipeline": {"bytes_per_frame": wl["bytes_per_frame"], "achieved": value / world * wl["bytes_per_frame"] / 1e9,
                                  "frac": value / world * wl["bytes_per_frame"] / 1e9 / peak},
                     "issue": issue_roofline(args.workload, value / world, clocks)},
        "stage_ms_per_step": stage,
        "status": status,
    }
    if match is not None:
        sm_hz = (clocks or {}).get("sm_mhz") or 1965.0
        # measured on this GPU (tools/pipe_rates.cu -> profiles/r02_pipe_rates.json): POPC 0.5, LOP3 2.0 warp instructions / clk / SM.
        # The carry-save Hamming distance costs 4 POPC + 16 LOP3 + 3 min/max per descriptor pair: 8 clk of POPC pipe, 9.5 clk of ALU pipe.
        match["popc_bound_gmatch_per_s"] = 148 * 0.5 * 32 * sm_hz * 1e6 / 4 / 1e9 * world
        match["alu_bound_gmatch_per_s"] = 148 * 2.0 * 32 * sm_hz * 1e6 / 19 / 1e9 * world
        match["popc_bound"] = ("148 SMs x 0.5 POPC warp-instructions/clk/SM (measured, profiles/r02_pipe_rates.json) x 32 lanes x SM clock / "
                               "4 POPC per pair (x N GPUs); alu_bound: 2.0 LOP3-class warp-instructions/clk/SM / 19 per pair")
        line["match"] = match
    if world == 1 and not args.no_latency:
        line["latency"] = latency_leg(X, wl, w, h, args.workload, cpu=not args.no_cpu_baseline and world == 1)
    if not args.no_cpu_baseline and world == 1:
        # the CPU arm runs in a FRESH process (this one holds pinned memory, a CUDA context and torch's thread pool, which made the
        # in-process sample up to 30 % slower than `--impl reference` on the same box): the reference arm itself, bounded sample
        def ref_arm(extra):
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--workload", args.workload] + extra,
                               capture_output=True, text=True, timeout=900)
            return json.loads(r.stdout.strip().splitlines()[-1])
        ref = ref_arm(["--steps", "3", "--warmup", "1"])
        line["cpu_baseline"] = dict(ref["cpu_baseline"])
        line["cpu_baseline"]["how"] = "bench.py --impl reference in a fresh subprocess (the same code path as the reference arm)"
        one = ref_arm(["--steps", "2", "--warmup", "1", "--cpu-threads", "1"])
        line["cpu_baseline"]["single_thread_frames_per_s"] = one["value"]
        line["cpu_baseline"]["single_thread_sample"] = one["cpu_baseline"]["sample"]
    emit(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


class StdoutToStderr:
    """Library banners (e.g. "NCCL version ..." printed on stdout by the NCCL runtime) must not mix with the ONE JSON line
    of the contract: file descriptor 1 points at stderr while the benchmark runs and is restored for the final print."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)
        return self

    def __exit__(self, *exc):
        sys.stdout.flush()
        os.dup2(self.saved, 1)
        os.close(self.saved)
        return False


def main():
    args = parse()
    lines = []
    with StdoutToStderr():
        if args.impl == "reference":
            run_reference(args, lines.append)
        else:
            run_ours(args, lines.append)
    for l in lines:
        print(l, flush=True)


if __name__ == "__main__":
    main()
