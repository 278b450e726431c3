#!/usr/bin/env python3
"""bench.py — frames/sec of ORB extract + Hamming match on synthetic frames (BASELINE.json metric).

A "step" is one pass of the hot path over one batch of B synthetic frames per GPU:
  extract (pyramid -> FAST cells -> octree -> orientation/blur/BRIEF) for every frame, then the
  frame-to-frame window match of SearchByProjection(Frame,Frame): keypoints of frame t-1 shifted by
  the known image motion, radius 15*scale[octave], octave band [o-1,o+1], 64x48 grid, first-min.
Default workload: configs[0]/C1 = 752x480, 1000 keypoints, 8 levels, scale 1.2 (the configuration
the metric is quoted on).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config C1|C3] [--batch B]
  python bench.py --impl reference ...     # the reference arm: CPU implementation on host cores

One JSON line on stdout (rank 0).  `value` = device-resident throughput, `e2e` = the same metric
through the host-buffer C-ABI call (H2D of frames + D2H of keypoints/descriptors/matches inside
the timed region), `roofline` for the dominant kernel, `cpu_baseline` = the oracle (CPU port of
the reference path) timed on this box's host cores.
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


def level_pixels(w, h, nlevels=8, scale=1.2):
    """Per-level pixel counts with the reference's float arithmetic (ORBextractor.cc:414-430,1174)."""
    sf = [np.float32(1.0)]
    for _ in range(1, nlevels):
        sf.append(np.float32(float(sf[-1]) * float(np.float32(scale))))
    out = []
    for s in sf:
        inv = np.float32(1.0) / s
        out.append(int(np.rint(np.float32(w) * inv)) * int(np.rint(np.float32(h) * inv)))
    return out


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


def cpu_reference(cfg, steps, warmup, sample_frames=None, budget_s=20.0):
    """The reference arm / cpu_baseline: the oracle's extract+match pipeline on all host threads."""
    import oracle
    oracle.build()
    w, h, nf, seed, _ = S.CONFIGS[cfg]
    lap = (0, 1000) if cfg == "C1" else (0, 0)
    threads = oracle.host_threads()
    n = sample_frames or max(2 * threads, 16)
    frames = S.make_frames(cfg, n)
    sh = shifts_for(0, n)
    run = lambda: oracle.bench_pipeline(frames, nf, ORB["scale"], ORB["nlevels"], ORB["ini"], ORB["mn"], lap, sh,
                                        MATCH_TH, threads)
    for _ in range(min(warmup, 1)):
        run()
    times, t_start = [], time.time()
    for _ in range(max(steps, 1)):
        sec, tk, tm = run()
        times.append(sec)
        if time.time() - t_start > budget_s:
            break
    sec = float(np.mean(times))
    return {"value": n / sec, "unit": "frames/s", "cores": threads, "kind": "port",
            "sample": f"{n} frames of {cfg} ({w}x{h}, {nf} kp) x {len(times)} passes, extract+match, "
                      f"{threads} threads, one frame per task", "ms_per_step": sec * 1e3, "steps": len(times)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C1", choices=["C1", "C3"])
    ap.add_argument("--batch", type=int, default=512, help="frames per step per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cfg = args.config
    w, h, nf, seed, _ = S.CONFIGS[cfg]
    lap = (0, 1000) if cfg == "C1" else (0, 0)
    workload = f"{cfg}: {w}x{h} mono, {nf} kp, 8 levels, scale 1.2, extract + frame-to-frame window match"

    if args.impl == "reference":
        if rank != 0:
            return
        cb = cpu_reference(cfg, args.steps, args.warmup)
        line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": "frames/s",
                "n_gpus": args.gpus, "steps": cb["steps"], "warmup": min(args.warmup, 1),
                "ms_per_step": cb["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u8", "data": "synthetic", "config": {"workload": workload},
                "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
                "e2e": {"value": cb["value"], "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return

    import torch
    import torch.distributed as dist
    from orb_slam3_vio_fixes_b200 import _lib
    from orb_slam3_vio_fixes_b200.extractor import ORBextractor
    from orb_slam3_vio_fixes_b200.matcher import ORBmatcher, results_to_numpy

    torch.cuda.set_device(local_rank)
    # keep this rank's host thread and its pinned staging buffers (first touch) on the NUMA node next to its GPU:
    # with N ranks the end-to-end leg is bound by host memory / PCIe root complexes, not by the kernels
    numa = "unset"
    all_cpus = os.sched_getaffinity(0)
    try:
        import pynvml
        pynvml.nvmlInit()
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local_rank))
        numa = "gpu-local"
    except Exception:
        pass
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B = args.batch
    t0 = rank * B  # every rank works on its own stretch of the sequence (weak scaling, no collective)
    frames_np = S.make_frames(cfg, B, t0=t0)
    pinned = torch.empty((B, h, w), dtype=torch.uint8, pin_memory=True)
    pinned.numpy()[:] = frames_np
    frames = pinned.to(dev)
    shift = torch.from_numpy(shifts_for(t0, B)).to(dev)
    src_frame = torch.from_numpy(((np.arange(B) - 1) % B).astype(np.int32)).to(dev)

    # `ex` runs the timed steps; `ex_t` records per-stage CUDA events (which switches the sub-batch
    # stream overlap off) and only serves the per-kernel breakdown below
    ex = ORBextractor(nf, ORB["scale"], ORB["nlevels"], ORB["ini"], ORB["mn"], max_batch=B, device=local_rank).plan(w, h)
    ex_t = ORBextractor(nf, ORB["scale"], ORB["nlevels"], ORB["ini"], ORB["mn"], max_batch=B, device=local_rank,
                        stage_timing=True).plan(w, h)
    sf = ex.GetScaleFactors()
    bounds = (0.0, float(w), 0.0, float(h))
    out = ex.alloc_outputs(B, dev)
    qbuf = None
    # resize x7, fast, octree, order, describe per sub-batch (orbx_extract_batch_device cuts batches >= 128 frames
    # in two, one per internal stream; ORBX_DEVICE_SPLIT) + grid, queries, search
    nsplit = int(os.environ.get("ORBX_DEVICE_SPLIT", "2"))
    nsplit = nsplit if B >= 64 * nsplit else 1
    launches_per_step = nsplit * ((ORB["nlevels"] - 1) + 4) + 3

    def step(src):
        nonlocal qbuf
        ex.extract_batch_device(src, lap, out)
        grid = ORBmatcher.grid_build(out["kps"], out["n"], bounds)
        qbuf = ORBmatcher.queries_shift(out["kps"], out["desc"], out["n"], src_frame, shift, MATCH_TH, sf, (-1, 1), qbuf)
        return ORBmatcher.search_area(out["kps"], out["desc"], out["n"], grid, bounds, qbuf["qx"], qbuf["qy"],
                                      qbuf["qr"], qbuf["qmin"], qbuf["qmax"], qbuf["qdesc"], qbuf["nq"])

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    # ---- device-resident throughput ("value") ---------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        res = step(frames)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        res = step(frames)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    # ---- per-kernel times over another K steps (CUDA events inside the library, same stream) ----
    stage_ms = {k: 0.0 for k in _lib.STAGES}
    m0, m1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    match_ms = 0.0
    for _ in range(args.steps):
        ex_t.extract_batch_device(frames, lap, out)
        m0.record()
        grid = ORBmatcher.grid_build(out["kps"], out["n"], bounds)
        qbuf = ORBmatcher.queries_shift(out["kps"], out["desc"], out["n"], src_frame, shift, MATCH_TH, sf, (-1, 1), qbuf)
        res = ORBmatcher.search_area(out["kps"], out["desc"], out["n"], grid, bounds, qbuf["qx"], qbuf["qy"],
                                     qbuf["qr"], qbuf["qmin"], qbuf["qmax"], qbuf["qdesc"], qbuf["nq"])
        m1.record()
        for k, v in ex_t.stage_times().items():
            stage_ms[k] += v / args.steps
        torch.cuda.synchronize()
        match_ms += m0.elapsed_time(m1) / args.steps
    clocks = sampler.stop() if rank == 0 else None
    tmax = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms = float(tmax.item())

    # ---- end to end through the host-buffer C-ABI call -------------------------------------------
    cap = ex.max_keypoints
    # Frames start in pinned HOST memory every step; the host-buffer C-ABI entry point copies them
    # to the device, extracts and copies {n, keypoints, descriptors} back; the match entry points
    # then score the device-resident copies and the match records are copied back — all inside the
    # timed region.
    res_pin = torch.empty((B, cap, 12), dtype=torch.uint8, pin_memory=True)
    kps_pin = torch.empty((B, cap, 7), dtype=torch.float32, pin_memory=True)
    desc_pin = torch.empty((B, cap, 32), dtype=torch.uint8, pin_memory=True)
    n_pin = torch.empty((B,), dtype=torch.int32, pin_memory=True)
    mono_pin = torch.empty((B,), dtype=torch.int32, pin_memory=True)
    host_out = (kps_pin.numpy(), desc_pin.numpy(), n_pin.numpy(), mono_pin.numpy())
    pinned_np = pinned.numpy()
    ex_host = ORBextractor(nf, ORB["scale"], ORB["nlevels"], ORB["ini"], ORB["mn"], max_batch=B,
                           device=local_rank).plan(w, h)
    dv = ex_host.last_outputs_device(B)
    qbuf2 = None

    def e2e():
        # orbx_extract_batch_host: H2D of the frames, the kernels and D2H of {n, mono, keypoints,
        # descriptors}, chunked and overlapped inside the C-ABI call (returns synchronised)
        nonlocal qbuf2
        ex_host.extract_batch_host(pinned_np, lap, out=host_out)
        grid = ORBmatcher.grid_build(dv["kps"], dv["n"], bounds)
        qbuf2 = ORBmatcher.queries_shift(dv["kps"], dv["desc"], dv["n"], src_frame, shift, MATCH_TH, sf, (-1, 1), qbuf2)
        r = ORBmatcher.search_area(dv["kps"], dv["desc"], dv["n"], grid, bounds, qbuf2["qx"], qbuf2["qy"], qbuf2["qr"],
                                   qbuf2["qmin"], qbuf2["qmax"], qbuf2["qdesc"], qbuf2["nq"])
        res_pin.copy_(r, non_blocking=True)
        torch.cuda.current_stream().synchronize()  # the step's results are on the host when it returns

    for _ in range(2):
        e2e()
    barrier()
    t_e2e = time.perf_counter()
    for _ in range(args.steps):
        e2e()
    barrier()
    e2e_s = time.perf_counter() - t_e2e
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s = float(te.item())
    h2d = B * w * h
    d2h = B * (4 + cap * 28 + cap * 32 + cap * 12)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline --------------------------------------------------------------------------------
    n_host = n_pin.numpy()
    kprime = float(n_host.mean())
    peaks = {}
    try:
        peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (of measured)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (of fallback)"
    resn = results_to_numpy(res_pin)
    matched = int(sum(((resn[f, :n_host[f]]["best_idx"] >= 0) & (resn[f, :n_host[f]]["best_dist"] <= 100)).sum()
                      for f in range(B)))
    sum_cand = 30.0 * kprime  # order of magnitude; exact count comes from tests
    ext_bytes, match_bytes = algorithmic_bytes(w, h, kprime, sum_cand, kprime, kprime)
    stages = []
    for k in _lib.STAGES:
        gbs = ext_bytes[k] * B / (stage_ms[k] * 1e-3) / 1e9 if stage_ms[k] > 0 else 0.0
        stages.append({"kernel": k, "ms": round(stage_ms[k], 4), "alg_bytes_per_frame": int(ext_bytes[k]),
                       "achieved_gbs": round(gbs, 1), "frac": round(gbs / peak, 4)})
    stages.append({"kernel": "match(grid+queries+search)", "ms": round(match_ms, 4),
                   "alg_bytes_per_frame": int(match_bytes),
                   "achieved_gbs": round(match_bytes * B / (match_ms * 1e-3) / 1e9, 1) if match_ms > 0 else 0.0})
    top = max(stages[:5], key=lambda s: s["ms"])
    # DRAM bytes of that kernel per launch from the committed `ncu --set full` capture (same command,
    # same batch), or null when no capture matches this batch size
    traffic = None
    try:
        caps = sorted((ROOT / "profiles").glob("*traffic_batch%d.json" % B))
        if caps:
            kern = json.loads(caps[-1].read_text())["kernels"].get("k_" + top["kernel"])
            if kern:
                # the capture's launches may cover a sub-batch (grid.y = frames in the launch): scale to the B
                # frames that `alg_bytes_per_launch` and the stage time refer to
                per = [(k_["dram_read_bytes"] + k_["dram_write_bytes"]) / max(1, int(k_["grid"].strip("()").split(",")[1]))
                       for k_ in kern]
                traffic = int(sum(per) / len(per) * B)
    except Exception:
        traffic = None
    total_bytes = sum(ext_bytes.values()) + match_bytes
    fps = world * B * args.steps / (ms * 1e-3)
    line = {
        "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": workload, "frames_per_step_per_gpu": B, "keypoints_per_frame": round(kprime, 1),
                   "matches_per_frame": round(matched / B, 1),
                   "l2_policy": f"inputs {B * w * h / 1e6:.0f} MB + intermediates exceed the 126 MB L2 every step",
                   "sharding": "independent frame batches per GPU, no collective", "host_affinity": numa},
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
    if not args.no_cpu_baseline:
        os.sched_setaffinity(0, all_cpus)  # the CPU baseline gets every host core again
        cb = cpu_reference(cfg, 3, 1, budget_s=15.0)
        line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
