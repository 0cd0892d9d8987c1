#!/usr/bin/env python3
"""bench.py -- headline benchmark of the ORB front end on B200 (contract: see DESIGN.md "Measurement").

One step = one pass of the hot path over a batch of synthetic KITTI-shaped stereo pairs
(BASELINE.json configs[1]): left+right ORB extraction (1241x376, 2000 features per camera, 8 levels,
1.2, FAST 20/7) followed by ComputeStereoMatches for every pair.  Metric: camera frames per second.

  python bench.py --gpus N --steps K --warmup W            # our CUDA path (N>1: torchrun, one rank per GPU)
  python bench.py --impl reference ...                     # the reference algorithm on the host cores

`value`  : device-resident (inputs already in HBM), CUDA events on the library's stream, max over ranks.
`e2e`    : same work through the public API from pinned host frames to host results.
`roofline`: dominant kernel, algorithmic bytes / CUDA-event time, vs MEASURED_PEAKS.json.
The oracle is used ONLY for the cpu_baseline leg and the reference arm.
"""
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

W, H, NFEAT = 1241, 376, 2000
MBF, FX = 386.1448, 718.856   # ref: ORB_SLAM2/Examples/Stereo/KITTI00-02.yaml:8,25
STAGES = ["h2d", "pyramid", "fast_cells", "octree", "describe", "d2h"]


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def make_pairs(n_distinct, seed0):
    from ar_orbslam2_b200.synth import synth_stereo_pair
    return [synth_stereo_pair(W, H, seed0 + i) for i in range(n_distinct)]


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}


def cpu_reference_run(pairs, n_threads, repeats=1):
    """The oracle (CPU restatement of the reference) on `pairs`, one pair per worker thread at a time;
    ctypes releases the GIL.  Returns (seconds, camera frames processed)."""
    from oracle import orbref
    mb = np.float32(MBF) / np.float32(FX)
    jobs = list(range(len(pairs))) * repeats
    lock = threading.Lock()
    it = iter(jobs)

    def worker():
        el, er = orbref.RefExtractor(NFEAT), orbref.RefExtractor(NFEAT)
        while True:
            with lock:
                j = next(it, None)
            if j is None:
                return
            l, r = pairs[j]
            kl, dl = el.extract(l)
            kr, dr = er.extract(r)
            orbref.stereo_match(el, er, kl, dl, kr, dr, MBF, mb)

    orbref.lib()
    ths = [threading.Thread(target=worker) for _ in range(n_threads)]
    t0 = time.perf_counter()
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    return time.perf_counter() - t0, 2 * len(jobs)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    sample_pairs = max(cores, 8)
    pairs = make_pairs(sample_pairs, 1000)
    for _ in range(args.warmup):
        cpu_reference_run(pairs[: max(1, cores // 2)], cores)
    t_total, frames = 0.0, 0
    for _ in range(args.steps):
        dt, n = cpu_reference_run(pairs, cores)
        t_total += dt
        frames += n
    v = frames / t_total
    sample = f"{sample_pairs} synthetic stereo pairs ({2 * sample_pairs} frames) per step, one pair per thread"
    line = {
        "impl": "reference", "metric": "orb_stereo_extract_frames_per_s", "value": v, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_total / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(sample_pairs),
        "cpu_baseline": {"value": v, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference cannot be compiled here (needs OpenCV 2.4 C++/Eigen/Pangolin); this is the oracle, the CPU "
                "restatement of ORBextractor.cc + Frame::ComputeStereoMatches, on all host cores",
    }
    print(json.dumps(line))


def workload_config(batch_pairs):
    return {"workload": "C2: KITTI-shaped stereo 1241x376, 2000 features/camera, 8 levels, scale 1.2, FAST 20/7: "
                        "left+right ORB extraction + ComputeStereoMatches",
            "stereo_pairs_per_step": batch_pairs, "frames_per_step": 2 * batch_pairs,
            "l2_policy": "inputs larger than L2 (frames/step x 466 KB > 126 MB) and a per-step working set > 1 GB",
            "parallelism": "independent frames sharded one batch per GPU, no collective"}


def run_ours(args):
    import torch
    import torch.distributed as dist

    from ar_orbslam2_b200 import _lib
    from ar_orbslam2_b200.extractor import ORBextractor
    from ar_orbslam2_b200.stereo import stereo_match_batch_device

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; liborbgpu has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    B = args.batch                       # stereo pairs per step (per GPU: weak scaling)
    NF = 2 * B
    pairs = make_pairs(args.distinct, 17 + 1000 * rank)
    host = torch.empty((NF, H, W), dtype=torch.uint8).pin_memory()
    hv = host.numpy()
    for b in range(B):
        l, r = pairs[b % len(pairs)]
        hv[b] = l
        hv[B + b] = r
    ex = ORBextractor(NFEAT, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=NF, device=local)
    cap = ex.default_capacity
    stream = torch.cuda.ExternalStream(ex.stream, device=dev)
    d_img = host.to(dev)
    d_kps = torch.zeros((NF, cap, 28), dtype=torch.uint8, device=dev)
    d_desc = torch.zeros((NF, cap, 32), dtype=torch.uint8, device=dev)
    d_n = torch.zeros(NF, dtype=torch.int32, device=dev)
    d_ur = torch.zeros((B, cap), dtype=torch.float32, device=dev)
    d_dp = torch.zeros((B, cap), dtype=torch.float32, device=dev)
    d_nm = torch.zeros(B, dtype=torch.int32, device=dev)
    h_n = torch.zeros(NF, dtype=torch.int32).pin_memory()
    h_kps = torch.zeros((NF, cap, 28), dtype=torch.uint8).pin_memory()
    h_desc = torch.zeros((NF, cap, 32), dtype=torch.uint8).pin_memory()
    h_ur = torch.zeros((B, cap), dtype=torch.float32).pin_memory()
    h_dp = torch.zeros((B, cap), dtype=torch.float32).pin_memory()
    h_nm = torch.zeros(B, dtype=torch.int32).pin_memory()
    mb = float(np.float32(MBF) / np.float32(FX))
    torch.cuda.synchronize()

    def step_device(src):
        ex.extract_batch_device(src.data_ptr(), NF, W, H, W, W * H, d_kps.data_ptr(), d_desc.data_ptr(), cap, d_n.data_ptr())
        stereo_match_batch_device(ex, 0, ex, B, B, d_kps.data_ptr(), d_desc.data_ptr(), d_n.data_ptr(),
                                  d_kps[B:].data_ptr(), d_desc[B:].data_ptr(), d_n[B:].data_ptr(), cap, MBF, mb,
                                  d_ur.data_ptr(), d_dp.data_ptr(), d_nm.data_ptr())

    d_in2 = torch.empty_like(d_img)

    def step_e2e():
        with torch.cuda.stream(stream):
            d_in2.copy_(host, non_blocking=True)
            step_device(d_in2)
            h_n.copy_(d_n, non_blocking=True)
            h_kps.copy_(d_kps, non_blocking=True)
            h_desc.copy_(d_desc, non_blocking=True)
            h_ur.copy_(d_ur, non_blocking=True)
            h_dp.copy_(d_dp, non_blocking=True)
            h_nm.copy_(d_nm, non_blocking=True)
        stream.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        s.record(stream)
        for _ in range(steps):
            fn()
        e.record(stream)
        e.synchronize()
        barrier()
        ms = s.elapsed_time(e)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    # ---- device-resident arm
    for _ in range(max(args.warmup, 3)):
        step_device(d_img)
    ex.check()
    n_kp = int(d_n.sum().item())
    n_matched = int(d_nm.sum().item())
    L = _lib.lib()
    L.orb_extractor_set_profiling(ex._h, 1)
    launches0 = ex.launch_count
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms = timed(lambda: step_device(d_img), args.steps)
    clocks = sampler.stop() if rank == 0 else None
    launches = ex.launch_count - launches0
    stage_ms = (C.c_double * 6)()
    calls = C.c_int64()
    L.orb_extractor_stage_times(ex._h, stage_ms, C.byref(calls))
    L.orb_extractor_set_profiling(ex._h, 0)
    ex.check()
    frames_total = NF * args.steps * world
    value = frames_total / (ms * 1e-3)

    # ---- end-to-end arm (pinned host frames in, host keypoints/descriptors/depth out)
    for _ in range(2):
        step_e2e()
    e2e_steps = max(3, args.steps // 2)
    barrier()
    t0 = time.perf_counter()
    ms_e2e = timed(step_e2e, e2e_steps)
    wall_e2e = time.perf_counter() - t0
    e2e_value = NF * e2e_steps * world / (ms_e2e * 1e-3)
    h2d = host.numel()
    d2h = sum(t.numel() * t.element_size() for t in (h_n, h_kps, h_desc, h_ur, h_dp, h_nm))
    assert int(h_n.sum().item()) == n_kp, "e2e arm produced different keypoint counts"

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (per-kernel algorithmic bytes, DESIGN.md "Kernels")
    hbm_peak, peak_src = load_peaks()
    per_call = [stage_ms[i] / max(calls.value, 1) for i in range(6)]
    P = W * H
    sizes = [ex.level_size(l) for l in range(8)]
    S_px = sum(w * h for w, h in sizes)
    win_px = sum((w - 38) * (h - 38) for w, h in sizes)
    kp_per_frame = n_kp / NF
    alg = {   # algorithmic bytes per frame, per kernel
        "pyramid": P + S_px,                       # read the source once, write the 8 levels once
        "fast_cells": S_px + 4 * 6 * kp_per_frame,  # read the 8 levels once, write ~6 candidates per kept keypoint
        "octree": 4 * 6 * kp_per_frame + 8 * kp_per_frame,
        "describe": 43 * 43 * kp_per_frame + 60 * kp_per_frame,
    }
    kern = {k: per_call[STAGES.index(k)] for k in alg}
    dom = max(kern, key=kern.get)
    achieved = alg[dom] * NF / (kern[dom] * 1e-3) / 1e9
    step_alg_bytes = (P * (1 + 2 * S_px / P) + 60 * kp_per_frame)   # SURVEY 8(d) per-frame figure
    step_gbs = step_alg_bytes * NF * args.steps / (ms * 1e-3) / 1e9

    # ---- CPU baseline: the oracle on the host cores, bounded sample
    cores = os.cpu_count() or 1
    sample_pairs = max(cores, 8)
    cpu_pairs = pairs[:sample_pairs] if len(pairs) >= sample_pairs else make_pairs(sample_pairs, 1000)
    reps = 1
    if args.skip_cpu:
        cpu_value = None
    else:
        cpu_reference_run(cpu_pairs[: max(1, cores // 2)], cores)
        dt, nfr = cpu_reference_run(cpu_pairs, cores, reps)
        while dt < 8.0 and reps < 64:
            reps *= 2
            dt, nfr = cpu_reference_run(cpu_pairs, cores, reps)
        cpu_value = nfr / dt

    line = {
        "metric": "orb_stereo_extract_frames_per_s", "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(B),
        "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e / e2e_steps, "wall_s": wall_e2e},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                     "frac": achieved / hbm_peak, "traffic": None, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": alg[dom] * NF, "kernel_ms": kern[dom]},
        "roofline_step": {"bound": "hbm", "achieved": step_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": step_gbs / hbm_peak,
                          "algorithmic_bytes_per_frame": step_alg_bytes, "formula": "P*(1+2S)+60N (SURVEY 8d)"},
        "stage_ms_per_step": dict({STAGES[i]: per_call[i] for i in range(6)},
                                  stereo_match=ms / args.steps - sum(per_call)),
        "keypoints_per_frame": kp_per_frame, "stereo_matches_per_pair": n_matched / B,
        "cpu_baseline": {"value": cpu_value, "unit": "frames/s", "cores": cores, "kind": "port",
                         "sample": f"{sample_pairs} synthetic stereo pairs x{reps} through the oracle, one pair per thread"},
    }
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=144, help="stereo pairs per step per GPU")
    ap.add_argument("--distinct", type=int, default=16, help="distinct synthetic pairs tiled into the batch")
    ap.add_argument("--skip-cpu", action="store_true", help="profiling runs: skip the CPU baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
