#!/usr/bin/env python3
"""bench.py — frames/sec of ORB extract + Hamming match on synthetic frames (BASELINE.json metric).

A "step" is one pass of the hot path over one batch of synthetic input per GPU:
  C1 / C3 / C5 (mono): extract (pyramid -> FAST strips -> octree -> orientation/blur/BRIEF) for B frames,
      then the frame-to-frame window match of SearchByProjection(Frame,Frame): keypoints of frame t-1 shifted
      by the known image motion, radius 15*scale[octave], octave band [o-1,o+1], 64x48 grid, first-min.
  C2 / C4 (stereo): extract both eyes of B pairs, then Frame::ComputeStereoMatches on the device
      (row-band Hamming -> 11x11 SAD sub-pixel refinement -> median cut); a "frame" is one stereo pair.
Default workload: C1 = 752x480, 1000 keypoints, 8 levels, scale 1.2 (the configuration the metric is quoted on).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config C1..C5] [--batch B]
  python bench.py --impl reference ...     # the reference arm: CPU implementation on the host cores

One JSON line on stdout (rank 0).  `value` = device-resident throughput, `e2e` = the same metric through the
host-buffer C-ABI call (H2D of the frames + D2H of keypoints/descriptors/matches inside the timed region, two
host threads with one handle each so that one call's tail overlaps the other's upload), `roofline` for the
dominant kernel, `cpu_baseline` = the faster of (1) the oracle (CPU port of the reference path) and (2) a cv2
arm (real OpenCV resize / FAST / GaussianBlur timed with T single-threaded processes + the oracle's
octree / orientation / BRIEF / match stage times), on this box's host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

from orb_slam3_vio_fixes_b200 import synthetic as S  # noqa: E402

METRIC = "frames/sec ORB extract+Hamming match @752x480,1k kp; % of HBM roofline"
ORB = dict(scale=1.2, nlevels=8, ini=20, mn=7)
MATCH_TH = 15.0  # SearchByProjection(Frame,Frame) mono window (Tracking.cc:2879-2894)
STEREO = {"C2": (458.654, 47.90639384423901), "C4": (718.856, 386.1448)}  # fx, bf (EuRoC / KITTI yaml)
DEFAULT_BATCH = {"C1": 512, "C2": 256, "C3": 512, "C4": 128, "C5": 32}
CPU_SAMPLE = {"C1": 256, "C2": 128, "C3": 256, "C4": 64, "C5": 16}  # frames (pairs) per CPU pass


def level_sizes(w, h, nlevels=8, scale=1.2):
    """Per-level (w, h) with the reference's float arithmetic (ORBextractor.cc:414-430,1174)."""
    sf = [np.float32(1.0)]
    for _ in range(1, nlevels):
        sf.append(np.float32(float(sf[-1]) * float(np.float32(scale))))
    out = []
    for s in sf:
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float32(w) * inv)), int(np.rint(np.float32(h) * inv))))
    return out


def level_pixels(w, h, nlevels=8, scale=1.2):
    return [a * b for a, b in level_sizes(w, h, nlevels, scale)]


def algorithmic_bytes(w, h, kprime, sum_cand, q, r):
    """per-frame algorithmic bytes, split by stage."""
    p = level_pixels(w, h)
    P = sum(p)
    ext = {
        "pyramid": 2 * P - p[0] - p[-1],          # read level l-1, write level l
        "fast": P,                                # FAST reads every level once
        "describe": 2 * P + kprime * (749 + 512) + kprime * 60,  # blur r+w, patch+samples, kp+desc out
        "octree": 0, "order": 0,                  # candidate lists / scratch are not counted
    }
    match = 32 * (q + r) + 4 * sum_cand + 12 * q
    return ext, match


def shifts_for(t0, n):
    """Image-plane shift that maps frame (f-1 mod n) keypoints into frame f, per frame."""
    out = np.zeros((n, 2), np.float32)
    for f in range(n):
        a, b = S.frame_shift(t0 + (f - 1) % n), S.frame_shift(t0 + f)
        out[f] = (a[0] - b[0], a[1] - b[1])
    return out


def workload_name(cfg):
    w, h, nf, _, stereo = S.CONFIGS[cfg]
    if stereo:
        return (f"{cfg}: {w}x{h} stereo pairs, {nf} kp/eye, 8 levels, scale 1.2, extract both eyes + "
                f"ComputeStereoMatches (row band + SAD refinement + median cut); a frame = one pair")
    return f"{cfg}: {w}x{h} mono, {nf} kp, 8 levels, scale 1.2, extract + frame-to-frame window match"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for nme, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nme)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------------
# CPU arms
# ------------------------------------------------------------------------------------------------------
_CV2_FRAMES = None
_CV2_SIZES = None


def _cv2_init():
    import cv2
    cv2.setNumThreads(1)


def _cv2_frame(i):
    """The OpenCV primitives ORBextractor runs per frame, by the real library: resize chain + copyMakeBorder
    (ComputePyramid), FAST at iniThFAST with NMS once per level (the reference calls it per 35-px cell: more calls,
    same pixels — this is the cheaper variant, i.e. a baseline in the CPU's favour), GaussianBlur 7x7 sigma 2."""
    import cv2
    img = _CV2_FRAMES[i]
    det = cv2.FastFeatureDetector_create(ORB["ini"], True)
    n = 0
    lvl = img
    for li, (lw, lh) in enumerate(_CV2_SIZES):
        if li:
            lvl = cv2.resize(lvl, (lw, lh), interpolation=cv2.INTER_LINEAR)
        cv2.copyMakeBorder(lvl, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
        n += len(det.detect(lvl[16:lh - 16, 16:lw - 16]))
        cv2.GaussianBlur(lvl, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
    return n


def cv2_arm_main(cfg, n, procs, passes):
    """Subprocess body of the cv2 arm (no CUDA in this process, so fork is safe): prints seconds per pass."""
    global _CV2_FRAMES, _CV2_SIZES
    import multiprocessing as mp
    w, h, nf, seed, stereo = S.CONFIGS[cfg]
    fr = S.make_frames(cfg, n)
    if stereo:
        fr = np.concatenate([fr, S.make_frames(cfg, n, right=True)])
    _CV2_FRAMES, _CV2_SIZES = fr, level_sizes(w, h)
    idx = list(range(fr.shape[0]))
    times = []
    with mp.get_context("fork").Pool(procs, initializer=_cv2_init) as pool:
        pool.map(_cv2_frame, idx[:max(procs, 8)])  # warm-up
        for _ in range(passes):
            t = time.perf_counter()
            pool.map(_cv2_frame, idx, chunksize=max(1, len(idx) // (4 * procs)))
            times.append(time.perf_counter() - t)
    print(json.dumps({"sec_per_pass": float(np.median(times)), "images": len(idx), "procs": procs}))


def cv2_arm(cfg, n, procs, passes=3):
    """Seconds per pass of the cv2 primitives over n frames (pairs) with `procs` single-threaded processes."""
    try:
        import cv2  # noqa: F401
    except Exception:
        return None
    try:
        out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--cv2-arm", cfg, str(n), str(procs), str(passes)],
                             capture_output=True, text=True, timeout=600)
        return json.loads(out.stdout.strip().splitlines()[-1])["sec_per_pass"]
    except Exception:
        return None


def cpu_port_mono(cfg, n, threads, warmup, passes, budget_s):
    import oracle
    w, h, nf, seed, _ = S.CONFIGS[cfg]
    lap = (0, 1000) if cfg == "C1" else (0, 0)
    frames = S.make_frames(cfg, n)
    sh = shifts_for(0, n)
    stages = np.zeros(8)
    acc = []

    def run():
        sec, tk, tm = oracle.bench_pipeline(frames, nf, ORB["scale"], ORB["nlevels"], ORB["ini"], ORB["mn"], lap, sh,
                                            MATCH_TH, threads, stages=stages)
        return sec

    for _ in range(warmup):
        run()
    t_start = time.time()
    for _ in range(max(passes, 1)):
        acc.append((run(), stages.copy()))
        if time.time() - t_start > budget_s:
            break
    sec = float(np.median([a[0] for a in acc]))
    st = np.median(np.stack([a[1] for a in acc]), axis=0)
    # what the cv2 arm does not replace: octree + orientation + descriptors (worker-seconds / threads) + match phase
    rest = float((st[2] + st[3] + st[5]) / threads + st[7])
    return sec, len(acc), rest


def cpu_port_stereo(cfg, n, threads, warmup, passes, budget_s):
    """Both eyes + ComputeStereoMatches per pair with the oracle, one pair per task on `threads` Python threads
    (every call is a ctypes call into C++, which releases the GIL)."""
    import oracle
    from concurrent.futures import ThreadPoolExecutor
    w, h, nf, seed, _ = S.CONFIGS[cfg]
    fx, bf = STEREO[cfg]
    mb, mbf = np.float32(bf / fx), np.float32(bf)
    left, right = S.make_frames(cfg, n), S.make_frames(cfg, n, right=True)
    tl = threading.local()
    sf = [np.float32(1.0)]
    for _ in range(1, ORB["nlevels"]):
        sf.append(np.float32(float(sf[-1]) * float(np.float32(ORB["scale"]))))
    sf = np.array(sf, np.float32)
    isf = (np.float32(1.0) / sf).astype(np.float32)

    def pair(i):
        if not hasattr(tl, "e"):  # one extractor per eye per worker, kept for the whole run
            tl.e = (oracle.Extractor(nf, ORB["scale"], ORB["nlevels"], ORB["ini"], ORB["mn"]),
                    oracle.Extractor(nf, ORB["scale"], ORB["nlevels"], ORB["ini"], ORB["mn"]))
        el, er = tl.e
        kl, dl, _ = el.extract(left[i], (0, 0))
        kr, dr, _ = er.extract(right[i], (0, 0))
        pl = [el.pyramid(l) for l in range(ORB["nlevels"])]
        pr = [er.pyramid(l) for l in range(ORB["nlevels"])]
        return oracle.stereo_matches(kl, dl, kr, dr, sf, isf, pl, pr, mb, mbf)[4]

    acc = []
    with ThreadPoolExecutor(threads) as pool:
        for _ in range(max(warmup, 1)):
            list(pool.map(pair, range(min(n, 2 * threads))))
        t_start = time.time()
        for _ in range(max(passes, 1)):
            t = time.perf_counter()
            list(pool.map(pair, range(n)))
            acc.append(time.perf_counter() - t)
            if time.time() - t_start > budget_s:
                break
    return float(np.median(acc)), len(acc), None


def cpu_reference(cfg, warmup, passes, budget_s=25.0, with_cv2=True, threads=None):
    """The reference arm / cpu_baseline: the reference path's CPU implementation on all host threads, on a bounded
    sample of the workload.  Two arms: the oracle port, and the cv2 arm (real OpenCV primitives for pyramid / FAST /
    blur, the port for the rest); the faster one is the baseline."""
    import oracle
    oracle.build()
    w, h, nf, seed, stereo = S.CONFIGS[cfg]
    threads = threads or oracle.host_threads()
    n = CPU_SAMPLE[cfg]
    eyes = 2 if stereo else 1
    if stereo:
        sec, npass, rest = cpu_port_stereo(cfg, n, threads, warmup, passes, budget_s)
    else:
        sec, npass, rest = cpu_port_mono(cfg, n, threads, warmup, passes, budget_s)
    port_fps = n / sec
    out = {"value": port_fps, "unit": "frames/s", "cores": threads, "kind": "port", "port_fps": round(port_fps, 1),
           "cv2_fps": None, "ms_per_step": sec * 1e3, "steps": npass}
    arm = "oracle port"
    if with_cv2 and rest is not None:
        c = cv2_arm(cfg, n, threads)
        if c is not None:
            cv2_fps = n / (c + rest)
            out["cv2_fps"] = round(cv2_fps, 1)
            if cv2_fps > port_fps:
                out["value"], arm = cv2_fps, "cv2 arm"
                out["ms_per_step"] = (c + rest) * 1e3
    out["sample"] = (f"{n} {'pairs' if stereo else 'frames'} of {cfg} ({w}x{h}, {nf} kp{'/eye' if stereo else ''}) x "
                     f"{npass} passes (median), extract+match, {threads} threads, one frame per task, extractors "
                     f"built outside the timed region; baseline = faster of oracle port ({out['port_fps']} fps) and "
                     f"cv2 arm ({out['cv2_fps']} fps: cv2 resize/FAST/GaussianBlur on {threads} single-threaded "
                     f"processes + the port's octree/orientation/BRIEF/match times) -> {arm}")
    return out


# ------------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------------
def sum_candidates(kps_q, kps_r, shift, sf, w, h):
    """Σ|cand(q)| of one frame pair, counted as GetFeaturesInArea does (Frame.cc:657-735): keypoints within
    |dx| < r and |dy| < r of the shifted query whose octave lies in [o-1, o+1]."""
    qx, qy = kps_q["x"] + shift[0], kps_q["y"] + shift[1]
    r = (MATCH_TH * sf[kps_q["octave"]]).astype(np.float32)
    dx = np.abs(kps_r["x"][None, :] - qx[:, None])
    dy = np.abs(kps_r["y"][None, :] - qy[:, None])
    lo, hi = kps_q["octave"][:, None] - 1, kps_q["octave"][:, None] + 1
    ok = (dx < r[:, None]) & (dy < r[:, None]) & (kps_r["octave"][None, :] >= lo) & (kps_r["octave"][None, :] <= hi)
    inside = (qx >= 0) & (qx < w) & (qy >= 0) & (qy < h)
    return int(ok[inside].sum())


def run_threads(workers, steps, device=None):
    """Runs `steps` calls spread round-robin over the worker callables, one host thread each; returns wall seconds."""
    errs = []

    def body(j):
        try:
            if device is not None:  # a new host thread starts on cuda:0: put it on this rank's GPU
                import torch
                torch.cuda.set_device(device)
            for _ in range(j, steps, len(workers)):
                workers[j]()
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    th = [threading.Thread(target=body, args=(j,)) for j in range(len(workers))]
    t = time.perf_counter()
    for x in th:
        x.start()
    for x in th:
        x.join()
    dt = time.perf_counter() - t
    if errs:
        raise errs[0]
    return dt


def latency_batch1(local_rank):
    """Median milliseconds of one-frame calls (latency-bound, PCIe + launch overheads): orbx_extract with host
    buffers, and the same followed by the eight orbx_get_pyramid(with_border=1) copies the compat operator()
    performs to refill mvImagePyramid (include/orbx_compat/ORBextractor.h)."""
    from orb_slam3_vio_fixes_b200.extractor import ORBextractor
    out = {}
    for cfg in ("C1", "C3"):
        w, h, nf, _, _ = S.CONFIGS[cfg]
        lap = (0, 1000) if cfg == "C1" else (0, 0)
        img = S.make_frames(cfg, 1)[0]
        ex = ORBextractor(nf, ORB["scale"], ORB["nlevels"], ORB["ini"], ORB["mn"], max_batch=1, device=local_rank).plan(w, h)
        for _ in range(5):
            ex(img, None, lap)
        t_ex, t_full = [], []
        for _ in range(40):
            t = time.perf_counter()
            ex(img, None, lap)
            t_ex.append(time.perf_counter() - t)
        for _ in range(20):
            t = time.perf_counter()
            ex(img, None, lap)
            for l in range(ORB["nlevels"]):
                ex.GetPyramid(l, with_border=True)
            t_full.append(time.perf_counter() - t)
        t_all = []
        for _ in range(20):
            t = time.perf_counter()
            ex(img, None, lap)
            ex.GetPyramidAll(with_border=True)
            t_all.append(time.perf_counter() - t)
        out[cfg] = {"orbx_extract_ms": round(float(np.median(t_ex)) * 1e3, 4),
                    "compat_operator_with_pyramid_copies_ms": round(float(np.median(t_all)) * 1e3, 4),
                    "with_eight_separate_pyramid_copies_ms": round(float(np.median(t_full)) * 1e3, 4)}
        ex.close()
    return out


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--cv2-arm":
        cv2_arm_main(sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5]))
        return
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C1", choices=["C1", "C2", "C3", "C4", "C5"])
    ap.add_argument("--batch", type=int, default=0, help="frames (pairs) per step per GPU")
    ap.add_argument("--e2e-threads", type=int, default=2, help="host threads (one handle each) of the end-to-end leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-latency", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cfg = args.config
    w, h, nf, seed, stereo = S.CONFIGS[cfg]
    lap = (0, 1000) if cfg == "C1" else (0, 0)
    workload = workload_name(cfg)
    warmup = max(args.warmup, 3)

    if args.impl == "reference":
        if rank != 0:
            return
        cb = cpu_reference(cfg, warmup, args.steps, budget_s=150.0)
        line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": "frames/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": warmup,
                "ms_per_step": cb["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u8", "data": "synthetic", "config": {"workload": workload},
                "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample", "port_fps", "cv2_fps")},
                "e2e": {"value": cb["value"], "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0, "passes_timed": cb["steps"]}
        print(json.dumps(line))
        return

    import torch
    import torch.distributed as dist
    from orb_slam3_vio_fixes_b200 import _lib
    from orb_slam3_vio_fixes_b200.extractor import ORBextractor
    from orb_slam3_vio_fixes_b200.matcher import ORBmatcher, results_to_numpy

    torch.cuda.set_device(local_rank)
    # keep this rank's host threads and its pinned staging buffers (first touch) on the NUMA node next to its GPU:
    # with N ranks the end-to-end leg is bound by host memory / PCIe root complexes, not by the kernels
    numa = "unset"
    all_cpus = os.sched_getaffinity(0)
    try:
        if os.environ.get("ORBX_BENCH_NO_AFFINITY"):
            raise RuntimeError("affinity off")
        import pynvml
        pynvml.nvmlInit()
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local_rank))
        numa = "gpu-local"
        # a VM may report a GPU-local set that is far smaller than the rank's fair share of the host (seen: a
        # handful of CPUs for 2 ranks on a 24-CPU box, e2e 30 k instead of 280 k frames/s): then keep every CPU
        if len(os.sched_getaffinity(0)) < max(4, len(all_cpus) // (2 * max(world, 1))):
            os.sched_setaffinity(0, all_cpus)
            numa = "all (gpu-local set too small)"
    except Exception:
        pass
    host_cpus = len(os.sched_getaffinity(0))
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B = args.batch or DEFAULT_BATCH[cfg]
    T = max(1, args.e2e_threads)
    # every rank works on its own stretch of the sequence / its own sequence (weak scaling, no collective):
    # C4 pair i -> rank i mod G, C5 sequence s -> rank s, C1-C3 consecutive stretches
    t0 = rank * B
    cap_pitch = (w + 15) & ~15  # device-resident frames keep 16-byte aligned rows (TMA / bulk-copy paths)

    def new_ex(timing=False):
        return ORBextractor(nf, ORB["scale"], ORB["nlevels"], ORB["ini"], ORB["mn"], max_batch=B, device=local_rank,
                            stage_timing=timing).plan(w, h)

    def pin_only(arr):
        pin = torch.empty(arr.shape, dtype=torch.uint8, pin_memory=True)
        pin.numpy()[:] = arr
        return pin

    def upload(arr):
        pin = pin_only(arr)
        d = torch.zeros((arr.shape[0], h, cap_pitch), dtype=torch.uint8, device=dev)
        d[:, :, :w] = pin.to(dev)
        return pin, d[:, :, :w]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    sf = None
    nsplit = int(os.environ.get("ORBX_DEVICE_SPLIT", "2"))
    nsplit = nsplit if B >= 64 * nsplit else 1
    extract_launches = nsplit * ((ORB["nlevels"] - 1) + 2 + 4)  # resize x7, fast x2, octree, order, blur|describe x2
    stage_ms = {k: 0.0 for k in _lib.STAGES}
    match_ms = 0.0
    m0, m1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    bounds = (0.0, float(w), 0.0, float(h))

    if not stereo:
        pinned, frames = upload(S.make_frames(cfg, B, t0=t0))
        shift_np = shifts_for(t0, B)
        shift = torch.from_numpy(shift_np).to(dev)
        src_frame = torch.from_numpy(((np.arange(B) - 1) % B).astype(np.int32)).to(dev)
        ex, ex_t = new_ex(), new_ex(True)
        sf = ex.GetScaleFactors()
        out = ex.alloc_outputs(B, dev)
        state = {"q": None, "grid": None, "res": None}
        launches_per_step = extract_launches + 3

        def match(o, st):
            st["grid"] = ORBmatcher.grid_build(o["kps"], o["n"], bounds, out=st["grid"])
            st["q"] = ORBmatcher.queries_shift(o["kps"], o["desc"], o["n"], src_frame, shift, MATCH_TH, sf, (-1, 1), st["q"])
            q = st["q"]
            st["res"] = ORBmatcher.search_area(o["kps"], o["desc"], o["n"], st["grid"], bounds, q["qx"], q["qy"], q["qr"],
                                               q["qmin"], q["qmax"], q["qdesc"], q["nq"], out=st["res"])
            return st["res"]

        def step():
            ex.extract_batch_device(frames, lap, out)
            return match(out, state)

        def step_timed():
            ex_t.extract_batch_device(frames, lap, out)
            m0.record()
            match(out, state)
            m1.record()
            return ex_t
    else:
        fx, bf = STEREO[cfg]
        mbf = np.float32(bf)
        max_d = mbf / np.float32(bf / fx)
        pin_l, left = upload(S.make_frames(cfg, B, t0=t0))
        pin_r, right = upload(S.make_frames(cfg, B, t0=t0, right=True))
        exl, exr, exl_t, exr_t = new_ex(), new_ex(), new_ex(True), new_ex(True)
        sf = exl.GetScaleFactors()
        ol, orr = exl.alloc_outputs(B, dev), exr.alloc_outputs(B, dev)
        launches_per_step = 2 * extract_launches + 3

        def stereo_match(a, b, o_l, o_r):
            bi, bd = ORBmatcher.stereo_rowband(o_l["kps"], o_l["desc"], o_l["n"], o_r["kps"], o_r["desc"], o_r["n"], sf, max_d)
            return ORBmatcher.stereo_refine(a, b, o_l["kps"], o_l["n"], o_r["kps"], bi, bd, mbf, max_d)

        def step():
            exl.extract_batch_device(left, (0, 0), ol)
            exr.extract_batch_device(right, (0, 0), orr)
            return stereo_match(exl, exr, ol, orr)

        def step_timed():
            exl_t.extract_batch_device(left, (0, 0), ol)
            exr_t.extract_batch_device(right, (0, 0), orr)
            m0.record()
            stereo_match(exl_t, exr_t, ol, orr)
            m1.record()
            return exl_t

    # ---- device-resident throughput ("value") ---------------------------------------------------
    for _ in range(warmup):
        res = step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        res = step()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    # ---- per-kernel times over another K steps (CUDA events inside the library, same stream) ----
    for _ in range(args.steps):
        et = step_timed()
        torch.cuda.synchronize()
        for k, v in et.stage_times().items():
            stage_ms[k] += v / args.steps
        match_ms += m0.elapsed_time(m1) / args.steps
    clocks = sampler.stop() if rank == 0 else None
    tmax = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms = float(tmax.item())

    # ---- end to end through the host-buffer C-ABI call -------------------------------------------
    # Every step's frames start in pinned HOST memory; orbx_extract_batch_host copies them to the device, extracts and
    # copies {n, keypoints, descriptors} back (chunked, overlapped, returns synchronised); the match entry points then
    # score the device-resident copies and the match records are copied back — all inside the timed region.  T host
    # threads with one handle (and one set of pinned buffers, one CUDA stream) each take the steps round-robin, the
    # way the reference runs its left / right extractors on two threads (Frame.cc:122-125): one call's compute / D2H
    # tail overlaps the other's upload.
    cap = (exl if stereo else ex).max_keypoints
    workers, keep = [], []
    for j in range(T):
        stream = torch.cuda.Stream(dev)
        if not stereo:
            fr_pin = pinned if j == 0 else pin_only(S.make_frames(cfg, B, t0=t0))
            exh = new_ex()
            host_out = tuple(torch.empty(s, dtype=d, pin_memory=True) for s, d in
                             (((B, cap, 7), torch.float32), ((B, cap, 32), torch.uint8), ((B,), torch.int32), ((B,), torch.int32)))
            res_pin = torch.empty((B, cap, 12), dtype=torch.uint8, pin_memory=True)
            dv = exh.last_outputs_device(B)
            st = {"q": None, "grid": None, "res": None}
            keep.append((exh, host_out, res_pin, fr_pin))

            def work(exh=exh, fr=fr_pin.numpy(), ho=tuple(t.numpy() for t in host_out), rp=res_pin, dv=dv, st=st, stream=stream):
                with torch.cuda.stream(stream):
                    exh.extract_batch_host(fr, lap, out=ho)
                    r = match(dv, st)
                    rp.copy_(r, non_blocking=True)
                    stream.synchronize()  # the step's results are on the host when it returns
        else:
            pl = pin_l if j == 0 else pin_only(S.make_frames(cfg, B, t0=t0))
            pr = pin_r if j == 0 else pin_only(S.make_frames(cfg, B, t0=t0, right=True))
            a, b = new_ex(), new_ex()
            hol = tuple(torch.empty(s, dtype=d, pin_memory=True) for s, d in
                        (((B, cap, 7), torch.float32), ((B, cap, 32), torch.uint8), ((B,), torch.int32), ((B,), torch.int32)))
            hor = tuple(torch.empty(s, dtype=d, pin_memory=True) for s, d in
                        (((B, cap, 7), torch.float32), ((B, cap, 32), torch.uint8), ((B,), torch.int32), ((B,), torch.int32)))
            ur_pin = torch.empty((B, cap), dtype=torch.float32, pin_memory=True)
            dp_pin = torch.empty((B, cap), dtype=torch.float32, pin_memory=True)
            dvl, dvr = a.last_outputs_device(B), b.last_outputs_device(B)
            keep.append((a, b, hol, hor, ur_pin, dp_pin, pl, pr))

            def work(a=a, b=b, pl=pl.numpy(), pr=pr.numpy(), hol=tuple(t.numpy() for t in hol),
                     hor=tuple(t.numpy() for t in hor), up=ur_pin, dp=dp_pin, dvl=dvl, dvr=dvr, stream=stream):
                with torch.cuda.stream(stream):
                    a.extract_batch_host(pl, (0, 0), out=hol)
                    b.extract_batch_host(pr, (0, 0), out=hor)
                    r = stereo_match(a, b, dvl, dvr)
                    up.copy_(r["u_right"], non_blocking=True)
                    dp.copy_(r["depth"], non_blocking=True)
                    stream.synchronize()
        workers.append(work)
    for wk in workers:
        wk()
        wk()
    barrier()
    e2e_s = run_threads(workers, args.steps, dev)
    if os.environ.get("ORBX_BENCH_VERBOSE"):
        print(f"[rank {rank}] e2e {args.steps} steps in {e2e_s * 1e3:.1f} ms", file=sys.stderr, flush=True)
    barrier()
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s = float(te.item())
    eyes = 2 if stereo else 1
    h2d = eyes * B * w * h
    d2h = eyes * B * (8 + cap * 28 + cap * 32) + (B * cap * 8 if stereo else B * cap * 12)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline --------------------------------------------------------------------------------
    peaks = {}
    try:
        peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (of measured)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (of fallback)"
    extra = {}
    if not stereo:
        n_host = keep[0][1][2].numpy()
        kprime = float(n_host.mean())
        resn = results_to_numpy(keep[0][2])
        matched = int(sum(((resn[f, :n_host[f]]["best_idx"] >= 0) & (resn[f, :n_host[f]]["best_dist"] <= 100)).sum()
                          for f in range(B)))
        kp_host = keep[0][1][0].numpy().view(np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                                                       ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")]))
        kp_host = kp_host.reshape(B, cap)
        sample = list(range(0, B, max(1, B // 8)))[:8]
        sc = [sum_candidates(kp_host[(f - 1) % B, :n_host[(f - 1) % B]], kp_host[f, :n_host[f]], shift_np[f],
                             np.asarray(sf, np.float32), w, h) for f in sample]
        sum_cand = float(np.mean(sc))
        ext_bytes, match_bytes = algorithmic_bytes(w, h, kprime, sum_cand, kprime, kprime)
        extra = {"keypoints_per_frame": round(kprime, 1), "matches_per_frame": round(matched / B, 1),
                 "sum_cand_per_frame": round(sum_cand, 1)}
    else:
        nl, nr = keep[0][2][2].numpy(), keep[0][3][2].numpy()
        kprime = float(nl.mean() + nr.mean()) / 2
        kept = float((keep[0][4].numpy() > 0).sum()) / B
        # row band: every left keypoint scores the right keypoints of its rows' band; Σ|cand| is not returned by the
        # kernel, so the 4·Σ|cand| index term is left out (descriptors + results only: a lower bound on the bytes)
        ext1, match_bytes = algorithmic_bytes(w, h, kprime, 0.0, float(nl.mean()), float(nr.mean()))
        ext_bytes = {k: 2 * v for k, v in ext1.items()}
        extra = {"keypoints_per_eye": round(kprime, 1), "stereo_matches_per_pair": round(kept, 1)}
    stages = []
    for k in _lib.STAGES:
        gbs = ext_bytes[k] * B / (stage_ms[k] * eyes * 1e-3) / 1e9 if stage_ms[k] > 0 else 0.0
        stages.append({"kernel": k, "ms": round(stage_ms[k] * eyes, 4), "alg_bytes_per_frame": int(ext_bytes[k]),
                       "achieved_gbs": round(gbs, 1), "frac": round(gbs / peak, 4)})
    stages.append({"kernel": "match", "ms": round(match_ms, 4), "alg_bytes_per_frame": int(match_bytes),
                   "achieved_gbs": round(match_bytes * B / (match_ms * 1e-3) / 1e9, 1) if match_ms > 0 else 0.0})
    top = max(stages[:5], key=lambda s: s["ms"])
    # DRAM bytes of that kernel per launch from the committed `ncu --set full` capture (same command, same batch),
    # or null when no capture matches this workload
    traffic = None
    try:
        caps = sorted((ROOT / "profiles").glob("*traffic_batch%d.json" % B)) if cfg == "C1" else []
        if caps:
            doc = json.loads(caps[-1].read_text())
            # the capture is ONE step of doc["batch"] frames (tools/profile_step.py B 1): every launch of the stage's
            # kernels in it, summed, is the stage's DRAM traffic for that many frames
            names = {"pyramid": ["k_resize_tile", "k_resize_walk", "k_resize"], "fast": ["k_fast_strip"],
                     "octree": ["k_octree"], "order": ["k_order"],
                     "describe": ["k_blur_levels", "k_describe_lite4", "k_describe_lite", "k_describe"]}[top["kernel"]]
            tot = sum(k_["dram_read_bytes"] + k_["dram_write_bytes"] for nm in names for k_ in doc["kernels"].get(nm, []))
            if tot > 0 and doc.get("steps_captured", 0) == 1:
                traffic = int(tot / doc["batch"] * B)
    except Exception:
        traffic = None
    total_bytes = sum(ext_bytes.values()) + match_bytes
    fps = world * B * args.steps / (ms * 1e-3)
    line = {
        "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": workload, "frames_per_step_per_gpu": B, **extra,
                   "l2_policy": f"inputs {eyes * B * w * h / 1e6:.0f} MB + intermediates exceed the 126 MB L2 every step",
                   "sharding": "independent frame batches per GPU, no collective", "host_affinity": numa, "host_cpus": host_cpus,
                   "e2e_host_threads": T},
        "e2e": {"value": world * B * args.steps / e2e_s, "unit": "frames/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h},
        "gpu_launches": launches_per_step * args.steps,
        "roofline": {"bound": "hbm", "kernel": top["kernel"], "achieved": top["achieved_gbs"], "peak": peak,
                     "unit": "GB/s", "frac": top["frac"], "traffic": traffic, "peak_source": peak_src,
                     "alg_bytes_per_launch": int(top["alg_bytes_per_frame"] * B)},
        "roofline_pipeline": {"alg_bytes_per_frame": int(total_bytes), "achieved_gbs": round(total_bytes * fps / world / 1e9, 1),
                              "frac": round(total_bytes * fps / world / 1e9 / peak, 4)},
        "stages": stages, "clocks": clocks,
    }
    if world == 1 and not args.no_latency:
        try:
            line["latency_batch1"] = latency_batch1(local_rank)
        except Exception as e:  # noqa: BLE001
            line["latency_batch1"] = {"error": str(e)}
    if not args.no_cpu_baseline:
        os.sched_setaffinity(0, all_cpus)  # the CPU baseline gets every host core again
        cb = cpu_reference(cfg, 1, 5, budget_s=20.0)
        line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample", "port_fps", "cv2_fps")}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
