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
import math
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


STAGE_KERNELS = {"pyramid": ("k_pyr_",), "fast_cells": ("k_fast_",), "octree": ("k_octree",), "describe": ("k_blur_", "k_describe"),
                 "stereo_match": ("k_stereo_",)}


def measured_stage_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per stage of one default step (288 frames of 1241x376) and the step's warp
    instruction count, read from the ncu launch list that profiles/CURRENT.json NAMES (`launch_shares`: a file written by
    tools/ncu_launch_list.py from an `ncu --metrics ...dram__bytes...` run of this same command; the entry is changed by hand in
    the commit that changes a kernel, so a stale list lying around under a later name is never picked up).
    Returns ({stage: bytes}, file, warp_instructions_per_288_frames)."""
    cur = os.path.join(ROOT, "profiles", "CURRENT.json")
    if not os.path.exists(cur):
        return {}, None, None
    name = json.load(open(cur)).get("launch_shares")
    path = os.path.join(ROOT, "profiles", name) if name else None
    if not path or not os.path.exists(path):
        return {}, None, None
    out = {k: 0.0 for k in STAGE_KERNELS}
    winst = 0.0
    for ln in open(path):
        f = ln.strip().split(",")
        if len(f) < 7 or not f[0].startswith("k_"):
            continue
        winst += float(f[4])
        for stage, prefixes in STAGE_KERNELS.items():
            if f[0].startswith(prefixes):
                out[stage] += (float(f[5]) + float(f[6])) * 1e6
    return out, os.path.relpath(path, ROOT), winst


_JSON_FD = None


def emit_json(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def make_pairs(n_distinct, seed0):
    from ar_orbslam2_b200.synth import synth_stereo_pair
    return [synth_stereo_pair(W, H, seed0 + i) for i in range(n_distinct)]


class SmiClockSampler:
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
        return {"sm_mhz": float(np.median(sm)), "sm_min_mhz": float(min(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


class ClockSampler:
    """SM clock and throttle reasons sampled every ~2 ms DURING the timed region through NVML (the timed region of the
    default run is ~30 ms: nvidia-smi's 100 ms loop sees it once or twice); falls back to the nvidia-smi loop."""

    def __init__(self, gpu_index):
        self.gpu, self.fallback, self.h, self.nv = gpu_index, None, None, None
        try:
            import pynvml as nv
            import torch
            nv.nvmlInit()
            try:
                uuid = str(torch.cuda.get_device_properties(gpu_index).uuid)
                self.h = nv.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            except Exception:
                vis = os.environ.get("CUDA_VISIBLE_DEVICES")
                idx = int(vis.split(",")[gpu_index]) if vis and vis.split(",")[gpu_index].isdigit() else gpu_index
                self.h = nv.nvmlDeviceGetHandleByIndex(idx)
            nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
            self.nv = nv
        except Exception:
            self.fallback = SmiClockSampler(gpu_index)
        self.sm, self.bits, self.run, self.th = [], 0, False, None

    def _loop(self):
        nv = self.nv
        while self.run:
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                self.bits |= int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        if self.fallback:
            return self.fallback.start()
        self.run = True
        self.th = threading.Thread(target=self._loop, daemon=True)
        self.th.start()

    def stop(self):
        if self.fallback:
            return self.fallback.stop()
        self.run = False
        self.th.join()
        nv = self.nv
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}
        try:
            mx = float(nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM))
        except Exception:
            mx = max(self.sm)
        return {"sm_mhz": float(np.median(self.sm)), "sm_min_mhz": float(min(self.sm)), "sm_max_mhz": mx, "reasons": sorted(n for b, n in names.items() if self.bits & b),
                "samples": len(self.sm), "source": "nvml, 2 ms period inside the timed region"}


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
    sample_pairs = 4 * max(cores, 8)
    pairs = make_pairs(min(sample_pairs, 32), 1000)
    pairs = [pairs[i % len(pairs)] for i in range(sample_pairs)]
    for _ in range(args.warmup):
        cpu_reference_run(pairs[: max(1, cores // 2)], cores)
    t_total, frames = 0.0, 0
    budget_s = 100.0   # the whole --steps K run stays within a couple of minutes whatever K the caller asks for
    for k in range(args.steps):
        dt, n = cpu_reference_run(pairs, cores)
        t_total += dt
        frames += n
        if k == 0 and args.steps * dt > budget_s:   # shrink the per-step sample (never below one pair per core)
            keep = max(cores, int(len(pairs) * budget_s / (args.steps * dt)))
            if keep < len(pairs):
                pairs = pairs[:keep]
                sample_pairs = f"{sample_pairs} (first step), then {keep}"
    v = frames / t_total
    # beside it, on ONE core (its driver's allocator is process-global): the reference's own ORBextractor.cc compiled unmodified
    # (oracle/_ref, OpenCV primitives = the cv2-pinned restatements) against the port on the same frames, extraction only
    ref_src = None
    try:
        from oracle import orbref
        if orbref.ref_available():
            fr = [pairs[i][0] for i in range(4)]
            a, b = orbref.ReferenceExtractor(NFEAT), orbref.RefExtractor(NFEAT)
            a.extract(fr[0]); b.extract(fr[0])
            t0 = time.perf_counter()
            for f in fr:
                a.extract(f)
            t1 = time.perf_counter()
            for f in fr:
                b.extract(f)
            t2 = time.perf_counter()
            ref_src = {"reference_source_frames_per_s_1core": len(fr) / (t1 - t0), "port_frames_per_s_1core": len(fr) / (t2 - t1),
                       "what": "extraction only (no stereo association), 4 frames of 1241x376, one thread each"}
    except Exception as exc:   # the checker's extra, never a reason to lose the line
        ref_src = {"unavailable": repr(exc)}
    sample = f"{sample_pairs} synthetic stereo pairs ({2 * sample_pairs} frames) per step, one pair per thread"
    line = {
        "impl": "reference", "metric": "orb_stereo_extract_frames_per_s", "value": v, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_total / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(args.batch),
        "cpu_baseline": {"value": v, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "reference_source": ref_src,
        "note": "the reference's build (cmake + OpenCV 2.4 C++ + Eigen + Pangolin) cannot run here; this is the oracle, the CPU "
                "restatement of ORBextractor.cc + Frame::ComputeStereoMatches, on all host cores; `reference_source` times the "
                "reference's own ORBextractor.cc (compiled against the cv2-pinned OpenCV stand-ins) beside the port on one core",
    }
    emit_json(line)


def workload_config(batch_pairs):
    return {"workload": "C2: KITTI-shaped stereo 1241x376, 2000 features/camera, 8 levels, scale 1.2, FAST 20/7: "
                        "left+right ORB extraction + ComputeStereoMatches",
            "stereo_pairs_per_step": batch_pairs, "frames_per_step": 2 * batch_pairs,
            "l2_policy": "inputs larger than L2 (frames/step x 466 KB > 126 MB) and a per-step working set > 1 GB",
            "parallelism": "independent frames sharded one batch per GPU, no collective"}


def golden_sweeps():
    path = os.path.join(ROOT, "tests", "golden", "oracle_sweep_hashes.json")
    return json.load(open(path)) if os.path.exists(path) else None


def host_copy_rate(torch, dev, host_t, dev_t, reps=3):
    """GB/s of the same pinned-host -> device copy the end-to-end arms issue (the ceiling of an H2D-bound step)."""
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dev_t.copy_(host_t, non_blocking=True)
    torch.cuda.synchronize()
    s.record()
    for _ in range(reps):
        dev_t.copy_(host_t, non_blocking=True)
    e.record()
    e.synchronize()
    return host_t.numel() * host_t.element_size() * reps / (s.elapsed_time(e) * 1e-3) / 1e9


def bench_c5(args, torch, dist, dev, local, rank, world, reduce_max, barrier):
    """Config C5 as BASELINE.json states it: brute-force top-2 of one frame's 2000 descriptors against a database of
    `--db-keyframes` keyframes x 2000 descriptors (50 000 x 2000 = 100 M rows = 3.2 GB), sharded by keyframe range over the
    ranks (STRONG scaling: the database is fixed, a rank holds K/N keyframes), local scan + NCCL all-gather of the partial
    top-2 (48 KB/rank) and of the per-keyframe accepted-match counts (4 B x K/N per rank) + device merge.  Gcompares/s is the
    whole-job rate; the roofline is the measured POPC issue rate (8 popc32 per compare).  N > 1: the merged result is compared
    with a one-GPU scan of the whole database on rank 0 and the run FAILS on any difference.  The CPU leg
    (__builtin_popcountll, all host cores) runs a 1/100 slice of the database and is extrapolated x100."""
    from ar_orbslam2_b200.matcher import DescriptorDB, measure_popc_rate
    from ar_orbslam2_b200.sharding import shard_keyframes, shard_range
    from ar_orbslam2_b200.synth import synth_descriptors, synth_queries
    K, KR, nq, reps, seed = args.db_keyframes, 2000, 2000, 3, 7
    TH_LOW, RATIO = 50, 0.75
    rb, re = shard_keyframes(K, KR, rank, world)
    db = DescriptorDB(rb, re, device=local)
    db.fill_synthetic(seed)
    nkf_local = (re - rb) // KR
    # queries: even rows are database rows (rank 0's shard) with 0..40 flipped bits, odd rows random
    q_host = torch.zeros((nq, 32), dtype=torch.uint8).pin_memory()
    if rank == 0:
        n0 = re - rb
        q_host.numpy()[:] = synth_queries(lambda idx: db.rows([int(i) % n0 for i in idx]), n0, nq, 3)
    q = q_host.to(dev)
    if world > 1:
        dist.broadcast(q, 0)
        q_host.copy_(q)
    st = torch.cuda.ExternalStream(db.stream, device=dev)
    i32, i64 = torch.int32, torch.int64
    loc = [torch.zeros(nq, dtype=dt, device=dev) for dt in (i32, i64, i32, i64)]
    out = [torch.zeros(nq, dtype=dt, device=dev) for dt in (i32, i64, i32, i64)]
    cnt_local = torch.zeros(nkf_local, dtype=i32, device=dev)
    width = max(e - b for b, e in (shard_range(K, r, world) for r in range(world)))
    cnt_pad = torch.zeros(width, dtype=i32, device=dev)
    cnt_all = torch.zeros(world * width, dtype=i32, device=dev)
    packed = torch.zeros((4, nq), dtype=i64, device=dev)
    gathered = torch.zeros((world, 4, nq), dtype=i64, device=dev)
    parts = [torch.zeros((world, nq), dtype=dt, device=dev) for dt in (i32, i64, i32, i64)]
    h_out = [torch.zeros(nq, dtype=dt).pin_memory() for dt in (i32, i64, i32, i64)]
    h_cnt = torch.zeros(world * width, dtype=i32).pin_memory()

    def one(q_dev):
        db.query_keyframes_device(q_dev.data_ptr(), nq, KR, TH_LOW, RATIO, *[t.data_ptr() for t in loc], cnt_local.data_ptr())
        if world == 1:
            return
        with torch.cuda.stream(st):
            for k in range(4):
                packed[k].copy_(loc[k])
            dist.all_gather_into_tensor(gathered, packed)
            for k in range(4):
                parts[k].copy_(gathered[:, k])   # preallocated: nothing is allocated on the library's stream
            db.merge_top2_device(world, nq, *[t.data_ptr() for t in parts], *[t.data_ptr() for t in out])
            cnt_pad[:nkf_local].copy_(cnt_local)
            dist.all_gather_into_tensor(cnt_all, cnt_pad)

    for _ in range(2):
        one(q)
    st.synchronize()
    l0 = db.launch_count
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    s.record(st)
    for _ in range(reps):
        one(q)
    e.record(st)
    e.synchronize()
    barrier()
    ms = reduce_max(s.elapsed_time(e))
    launches = int(db.launch_count - l0)
    res = out if world > 1 else loc

    def counts_now():
        if world == 1:
            return cnt_local.cpu().numpy()
        a = cnt_all.cpu().numpy()
        owned = [shard_range(K, rr, world) for rr in range(world)]
        return np.concatenate([a[rr * width: rr * width + (ke - kb)] for rr, (kb, ke) in enumerate(owned)])

    # end to end: the queries come from pinned host memory and the merged top-2 + all keyframe counts go back to it
    s2, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    q2 = torch.empty_like(q)
    barrier()
    with torch.cuda.stream(st):
        s2.record(st)
        for _ in range(reps):
            q2.copy_(q_host, non_blocking=True)
            one(q2)
            for k in range(4):
                h_out[k].copy_(res[k], non_blocking=True)
            if world == 1:
                h_cnt[:nkf_local].copy_(cnt_local, non_blocking=True)
            else:
                h_cnt.copy_(cnt_all, non_blocking=True)
        e2.record(st)
    e2.synchronize()
    barrier()
    ms_e2e = reduce_max(s2.elapsed_time(e2))
    total_rows = K * KR
    gcmp = total_rows * nq * reps / (ms * 1e-3) / 1e9
    popc = measure_popc_rate()
    counts = counts_now()
    r = {"workload": f"C5: {K} keyframes x {KR} descriptors = {total_rows} rows ({total_rows * 32 / 1e9:.2f} GB), {nq} queries, "
                     f"sharded by keyframe range over {world} GPU(s)",
         "scaling": "strong", "gcompares_per_s": gcmp, "queries": nq, "db_rows_total": total_rows, "db_rows_per_gpu": (re - rb),
         "ms_per_query_batch": ms / reps,
         "e2e": {"gcompares_per_s": total_rows * nq * reps / (ms_e2e * 1e-3) / 1e9, "ms_per_query_batch": ms_e2e / reps,
                 "h2d_bytes_per_step": nq * 32, "d2h_bytes_per_step": nq * 24 + 4 * K},
         "roofline": {"bound": "int/popc issue", "achieved": gcmp * 8 / world, "peak": popc, "unit": "Gpopc32/s per GPU",
                      "frac": gcmp * 8 / (popc * world), "peak_source": "measured in this run (orb_measure_popc_rate)",
                      "note": "HBM traffic is 32 B per row per 256 queries: 0.4 % of the HBM roofline, the scan is POPC-issue bound"},
         "launches": launches,
         "exchange": "none" if world == 1 else "nccl all_gather of 48 KB/rank (top-2) + 4 B x K/N per rank (keyframe counts) + device merge",
         "planted_queries_found": bool((res[0][0::2] <= 40).all().item()),
         "keyframe_counts": {"rule": "best <= 50 and best < 0.75 * second inside the keyframe (ORBmatcher.cc:216-231)",
                             "keyframes_with_matches": int((counts > 0).sum()), "accepted_pairs": int(counts.sum()),
                             "max_per_keyframe": int(counts.max(initial=0))}}
    if world > 1 and rank == 0:   # the same rows on ONE GPU: the merged result must be identical
        full = DescriptorDB(0, total_rows, device=local)
        full.fill_synthetic(seed)
        d1, i1, d2, i2, c1 = full.query_keyframes(q_host.numpy(), KR, TH_LOW, RATIO)
        full.close()
        same = (np.array_equal(d1, out[0].cpu().numpy()) and np.array_equal(i1, out[1].cpu().numpy()) and
                np.array_equal(d2, out[2].cpu().numpy()) and np.array_equal(i2, out[3].cpu().numpy()) and np.array_equal(c1, counts))
        r["merged_equals_one_gpu_scan"] = bool(same)
        if not same:
            raise SystemExit("bench.py: the sharded top-2 / keyframe counts differ from a one-GPU scan of the same database")
    if rank == 0 and not args.skip_cpu:
        # CPU leg on a 1/100 slice (whole keyframes), and the GPU against it on the same slice
        from oracle import orbref
        cores = os.cpu_count() or 1
        ks = max(1, K // 100)
        rows = synth_descriptors(ks * KR, seed)
        qh = q_host.numpy()
        t0 = time.perf_counter()
        want = orbref.hamming_top2(qh, rows, fast=True, threads=cores)
        t1 = time.perf_counter()
        wc = orbref.db_keyframe_counts(qh, rows, KR, TH_LOW, RATIO, threads=cores)
        t2 = time.perf_counter()
        sl = DescriptorDB(0, ks * KR, device=local)
        sl.fill_synthetic(seed)
        got = sl.query_keyframes(qh, KR, TH_LOW, RATIO)
        sl.close()
        r["cpu_baseline"] = {"value": ks * KR * nq / (t1 - t0) / 1e9, "unit": "Gcompares/s", "cores": cores, "kind": "port",
                             "sample": f"1/{K // ks} slice ({ks} keyframes = {ks * KR} rows) x {nq} queries, __builtin_popcountll top-2 scan; "
                                       f"whole database extrapolated: {(t1 - t0) * K / ks:.1f} s per query batch",
                             "keyframe_counts_gcompares_per_s": ks * KR * nq / (t2 - t1) / 1e9}
        r["slice_equals_oracle"] = bool(all(np.array_equal(a, b) for a, b in zip(got[:4], want)) and np.array_equal(got[4], wc))
        if not r["slice_equals_oracle"]:
            raise SystemExit("bench.py: GPU top-2 / keyframe counts differ from the CPU oracle on the 1/100 slice")
    _KEEP.append(db)   # its stream must outlive the pinned buffers that were copied on it (see the note in run_ours)
    return r, popc


def bench_frames_leg(args, torch, dist, dev, local, rank, world, reduce_max, barrier, *, w, h, nfeat, total, chunk, steps,
                     golden_key, label, stages=False, cpu_seconds=5.0):
    """`total` independent frames of w x h, total / N per rank (config C4: 256 frames of 1920x1080; config C1: 640x480), no
    collective.  Two arms: device-resident (one batched call per step) and end to end FROM HOST MEMORY (pinned frames ->
    device in chunks of `chunk` frames on a copy stream while the previous chunk computes -> results back to pinned host
    memory per chunk).  Every frame of the end-to-end arm is checked by SHA-256 against the oracle's hashes
    (tests/golden/oracle_sweep_hashes.json); the CPU oracle runs a bounded sample on all host cores on rank 0."""
    from ar_orbslam2_b200 import _lib
    from ar_orbslam2_b200.extractor import ORBextractor
    from ar_orbslam2_b200.hashing import frame_hash
    from ar_orbslam2_b200.synth import synth_frame
    gold = (golden_sweeps() or {}).get(golden_key)
    nseeds = len(gold) if gold else 16
    assert total % world == 0, "frames must divide over the ranks"
    nb, b0 = total // world, rank * (total // world)
    seeds = [(b0 + i) % nseeds for i in range(nb)]
    distinct = {sd: synth_frame(w, h, sd) for sd in sorted(set(seeds))}
    host = torch.empty((nb, h, w), dtype=torch.uint8).pin_memory()
    hv = host.numpy()
    for i, sd in enumerate(seeds):
        hv[i] = distinct[sd]
    ex = ORBextractor(nfeat, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=nb, device=local)
    _KEEP.append(ex)
    cap = ex.default_capacity
    d_k = torch.zeros((nb, cap, 28), dtype=torch.uint8, device=dev)
    d_d = torch.zeros((nb, cap, 32), dtype=torch.uint8, device=dev)
    d_n = torch.zeros(nb, dtype=torch.int32, device=dev)
    h_k, h_d, h_n = (torch.zeros(tuple(t.shape), dtype=t.dtype).pin_memory() for t in (d_k, d_d, d_n))
    cs = torch.cuda.ExternalStream(ex.stream, device=dev)
    d_img = host.to(dev)
    run = lambda: ex.extract_batch_device(d_img.data_ptr(), nb, w, h, w, w * h, d_k.data_ptr(), d_d.data_ptr(), cap, d_n.data_ptr())
    for _ in range(3):
        run()
    ex.check()
    l0 = ex.launch_count
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    s.record(cs)
    for _ in range(steps):
        run()
    e.record(cs)
    e.synchronize()
    barrier()
    ms = reduce_max(s.elapsed_time(e))
    launches = int(ex.launch_count - l0)
    n_kp = float(d_n.sum().item()) / nb
    sizes = [ex.level_size(l) for l in range(8)]
    S_px = sum(a * b for a, b in sizes)
    alg_frame = w * h + 2 * S_px + 60 * n_kp            # P(1 + 2S) + 60 N
    hbm_peak, peak_src = load_peaks()
    r = {"workload": label, "frames_total": total, "frames_per_gpu": nb, "scaling": "strong" if world > 1 else "single GPU",
         "frames_per_s": total * steps / (ms * 1e-3), "ms_per_batch": ms / steps, "keypoints_per_frame": n_kp, "launches": launches,
         "roofline_step": {"bound": "hbm", "achieved": alg_frame * total * steps / (ms * 1e-3) / 1e9 / world, "peak": hbm_peak,
                           "unit": "GB/s per GPU", "frac": alg_frame * total * steps / (ms * 1e-3) / 1e9 / world / hbm_peak,
                           "algorithmic_bytes_per_frame": alg_frame, "formula": "P*(1+2S)+60N (SURVEY 8d)", "peak_source": peak_src}}
    if stages and rank == 0:   # serial schedule with the library's stage events: where a batch of large frames spends its time
        L = _lib.lib()
        L.orb_extractor_set_profiling(ex._h, 1)
        for _ in range(steps):
            run()
        ex.sync()
        sm = (C.c_double * 6)()
        calls = C.c_int64()
        L.orb_extractor_stage_times(ex._h, sm, C.byref(calls))
        L.orb_extractor_set_profiling(ex._h, 0)
        r["stage_ms_per_batch_serial"] = {STAGES[i]: sm[i] / max(calls.value, 1) for i in range(6)}
    # ---- end to end from host memory, chunked: at least four chunks per rank whatever the share, so that uploads, kernels and
    # downloads overlap even when total / N is as small as the nominal chunk (C4 at N = 8: 32 frames per rank)
    chunk = max(1, min(chunk, (nb + 3) // 4))
    nch = (nb + chunk - 1) // chunk
    sin, sout = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    d_in = [torch.empty((min(chunk, nb), h, w), dtype=torch.uint8, device=dev) for _ in range(2)]
    ev = lambda n: [torch.cuda.Event() for _ in range(n)]
    in_ready, in_free, out_ready, out_free = ev(2), ev(2), ev(nch), ev(nch)
    used_in, used_out = [False, False], [False] * nch

    def e2e_step():
        for c in range(nch):
            c0, c1, sl = c * chunk, min((c + 1) * chunk, nb), c % 2
            if used_in[sl]:
                sin.wait_event(in_free[sl])
            with torch.cuda.stream(sin):
                d_in[sl][: c1 - c0].copy_(host[c0:c1], non_blocking=True)
                in_ready[sl].record(sin)
            cs.wait_event(in_ready[sl])
            if used_out[c]:
                cs.wait_event(out_free[c])         # the previous step's results of this chunk have left the device arrays
            ex.extract_batch_device(d_in[sl].data_ptr(), c1 - c0, w, h, w, w * h, d_k[c0:].data_ptr(), d_d[c0:].data_ptr(), cap,
                                    d_n[c0:].data_ptr())
            in_free[sl].record(cs)
            out_ready[c].record(cs)
            sout.wait_event(out_ready[c])
            with torch.cuda.stream(sout):
                h_k[c0:c1].copy_(d_k[c0:c1], non_blocking=True)
                h_d[c0:c1].copy_(d_d[c0:c1], non_blocking=True)
                h_n[c0:c1].copy_(d_n[c0:c1], non_blocking=True)
                out_free[c].record(sout)
            used_in[sl] = used_out[c] = True

    e2e_step()
    torch.cuda.synchronize()
    s2, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    s2.record(sin)
    cs.wait_event(s2)
    for _ in range(steps):
        e2e_step()
    sout.wait_stream(cs)
    e2.record(sout)
    e2.synchronize()
    barrier()
    ms_e2e = reduce_max(s2.elapsed_time(e2))
    ex.check()
    link = host_copy_rate(torch, dev, host, d_img)
    d2h = sum(t.numel() * t.element_size() for t in (h_k, h_d, h_n))
    r["e2e"] = {"frames_per_s": total * steps / (ms_e2e * 1e-3), "ms_per_batch": ms_e2e / steps, "h2d_bytes_per_step": host.numel(),
                "d2h_bytes_per_step": d2h, "chunk_frames": chunk, "host_link_GBps": link,
                "copy_only_frames_per_s": world * link * 1e9 / (w * h),
                "pipeline": f"pinned host frames -> device in chunks of {chunk} on a copy stream while the previous chunk computes; "
                            "results -> pinned host per chunk on a second copy stream"}
    if gold:
        n, K_, D_ = h_n.numpy(), h_k.numpy(), h_d.numpy()
        bad = [i for i in range(nb) if int(n[i]) != gold[seeds[i]]["n"] or frame_hash(K_[i], D_[i], int(n[i])) != gold[seeds[i]]["sha256"]]
        ok = torch.tensor([0 if bad else 1], device=dev)
        if world > 1:
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        r["all_frames_equal_oracle_sha256"] = bool(ok.item())
        if bad:
            raise SystemExit(f"bench.py: {golden_key} frames {bad[:8]} of rank {rank} differ from the oracle's hashes")
    if rank == 0 and not args.skip_cpu:
        from oracle import orbref
        cores = os.cpu_count() or 1
        frames = [distinct[sd] for sd in sorted(distinct)]
        one = orbref.RefExtractor(nfeat)
        t1 = []
        for f in frames[:8]:
            t0 = time.perf_counter(); one.extract(f); t1.append(time.perf_counter() - t0)
        lock, done = threading.Lock(), [0]
        t_end = time.perf_counter() + cpu_seconds

        def worker(tid):
            exr, i = orbref.RefExtractor(nfeat), tid
            while time.perf_counter() < t_end:
                exr.extract(frames[i % len(frames)])
                i += cores
                with lock:
                    done[0] += 1

        ths = [threading.Thread(target=worker, args=(t,)) for t in range(cores)]
        t0 = time.perf_counter()
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        dt = time.perf_counter() - t0
        r["cpu_baseline"] = {"value": done[0] / dt, "unit": "frames/s", "cores": cores, "kind": "port",
                             "sample": f"{done[0]} extractions of the leg's frames in {dt:.1f} s, one RefExtractor per thread",
                             "one_thread_ms_median": 1e3 * float(np.median(t1)),
                             "one_thread_pattern": "per-frame wall time around the extractor call, median (Examples/Monocular/mono_tum.cc:79-88)"}
    return r


def bench_c3(args, popc_rate=None):
    """Config C3 (BASELINE.json configs[2]): 752x480, frame B = frame A shifted by (+7,+4), init extractor (2000
    features).  Per-call latency of the three tracking matchers through the C ABI from host arrays, measured by a C++ caller
    (ar_orbslam2_b200/shim/c3_bench: steady_clock around each call, median of `reps`), beside the reference's own
    ORBmatcher.cc (oracle/_ref/liborbmatcher_ref.so, compiled unmodified) on one host core with a timer around the same
    method call, on the same inputs; results compared array for array."""
    import tempfile
    from ar_orbslam2_b200.extractor import ORBextractor
    from ar_orbslam2_b200.synth import SplitMix64, shifted_frame, synth_frame
    W3, H3, reps = 752, 480, 50
    FX3 = FY3 = np.float32(435.2); CX3, CY3 = np.float32(367.2), np.float32(252.2)   # ref Examples/Stereo/EuRoC.yaml:8-11
    a = synth_frame(W3, H3, 31)
    b = shifted_frame(a, 7, 4, 32)
    ex = ORBextractor(2000, 1.2, 8, 20, 7, max_width=W3, max_height=H3)
    ka, da = ex(a)
    kb, db = ex(b)
    scale = ex.GetScaleFactors()
    ex.close()
    n = len(ka)
    # map points of the Last frame: depth z = 2 + (hash(i) mod 1000)/100 m, Last pose I, Cur pose translated by (0.05, 0.02, 0) m
    z = np.array([2 + (SplitMix64(i).next() % 1000) / 100 for i in range(n)], np.float32)
    xyz = np.stack([(ka["x"] - CX3) * z / FX3, (ka["y"] - CY3) * z / FY3, z], 1).astype(np.float32)
    has = (np.arange(n) % 17 != 0).astype(np.uint8)
    obs = np.ones(n, np.uint8)
    claimed_last = np.zeros(n, np.uint8)
    Tl = np.eye(4, dtype=np.float32)
    Tc = np.eye(4, dtype=np.float32); Tc[:3, 3] = (0.05, 0.02, 0.0)
    # local map points for SearchByProjection(Frame, MapPoints): A's features projected into B
    rng = SplitMix64(9)
    in_view = np.array([rng.next() % 10 != 0 for _ in range(n)], np.uint8)
    px, py = (ka["x"] + 7).astype(np.float32), (ka["y"] + 4).astype(np.float32)
    pxr = (px - 6).astype(np.float32)
    lvl = ka["octave"].astype(np.int32)
    vc = np.full(n, 0.9, np.float32)
    claimed_pts = np.zeros(len(kb), np.uint8)
    out = {"workload": f"C3: {W3}x{H3}, frame B = A shifted (+7,+4), 2000-feature init extractor: {n} / {len(kb)} key points",
           "how": "GPU: C++ caller, steady_clock around each C-ABI call from host arrays, frame indices cached (as the C++ shim does); "
                  "reference: the reference's own ORBmatcher.cc compiled unmodified, timer around the method call, one core"}
    # ---- reference (CPU) side
    ref = None
    try:
        from oracle import orbref
        from ar_orbslam2_b200.tracking import Frame, ORBmatcher
        if orbref.ref_matcher_available():
            rb = orbref.ReferenceSourceBackend()
            rb.L.refm_last_call_us.restype = C.c_double
            bounds = (0.0, float(W3), 0.0, float(H3))
            mk = lambda k, d: Frame(k, d, scale, bounds, backend=rb)
            r1, r2 = mk(ka, da), mk(kb, db)
            prev0 = np.stack([ka["x"], ka["y"]], 1).astype(np.float32)
            t = []
            for _ in range(12):
                pr = prev0.copy()
                nm_i, m12_ref = ORBmatcher(0.9, True, backend=rb).SearchForInitialization(r1, r2, pr, 100)
                t.append(rb.L.refm_last_call_us())
            ref = {"search_for_initialization": {"us_median": float(np.median(t)), "nmatches": nm_i}}
            prev_ref = pr
            rc, rl = mk(ka, da), mk(ka, da)
            t = []
            for _ in range(12):
                R = orbref.refm_search_by_projection_last(rc, rl, xyz, has, obs, claimed_last, Tc, Tl, FX3, FY3, CX3, CY3, 40.0, 0.11, 15.0, True,
                                                          0.9, True)
                t.append(rb.L.refm_last_call_us())
            ref["search_by_projection_last"] = {"us_median": float(np.median(t)), "nmatches": R["nmatches"]}
            t = []
            for _ in range(12):
                nm_p, asg_p_ref = ORBmatcher(0.8, backend=rb).SearchByProjectionPoints(r2, in_view, px, py, pxr, lvl, vc, da, 3.0, claimed_pts, obs)
                t.append(rb.L.refm_last_call_us())
            ref["search_by_projection_points"] = {"us_median": float(np.median(t)), "nmatches": nm_p}
    except Exception as exc:   # the checker's leg must not cost the line
        ref = {"unavailable": repr(exc)}
    if not ref or "unavailable" in ref:
        out["reference"] = ref or {"unavailable": "oracle/_ref/liborbmatcher_ref.so not built"}
        return out
    # ---- GPU side: the C++ caller
    exe = os.path.join(ROOT, "ar_orbslam2_b200", "shim", "c3_bench")
    with tempfile.TemporaryDirectory() as td:
        w = lambda name, arr: np.ascontiguousarray(arr).tofile(os.path.join(td, name))
        w("a.raw", a); w("b.raw", b); w("ka.bin", ka); w("da.bin", da); w("kb.bin", kb); w("db.bin", db); w("scale.bin", scale.astype(np.float32))
        w("last_valid.bin", R["valid"]); w("last_u.bin", R["u"]); w("last_v.bin", R["v"]); w("last_invz.bin", R["invz"])
        w("last_octave.bin", ka["octave"].astype(np.int32)); w("last_angle.bin", ka["angle"].astype(np.float32)); w("last_obs.bin", obs)
        w("last_claimed.bin", claimed_last); w("last_uright.bin", np.full(n, -1, np.float32))
        w("pts_in_view.bin", in_view); w("pts_x.bin", px); w("pts_y.bin", py); w("pts_xr.bin", pxr); w("pts_level.bin", lvl)
        w("pts_view_cos.bin", vc); w("pts_obs.bin", obs); w("pts_claimed.bin", claimed_pts)
        p = subprocess.run([exe, td, str(W3), str(H3), str(reps)], capture_output=True, text=True)
        if p.returncode != 0:
            out["gpu"] = {"failed": p.stderr[-400:]}
            return out
        g = json.loads(p.stdout.strip().splitlines()[-1])
        rd = lambda name, dt: np.fromfile(os.path.join(td, name), dt)
        same = {"search_for_initialization": bool(np.array_equal(rd("init_m12.bin", np.int32), m12_ref) and
                                                  np.array_equal(rd("init_prev.bin", np.float32).reshape(-1, 2), prev_ref)),
                "search_by_projection_last": bool(np.array_equal(np.where(rd("last_assigned.bin", np.int32) < 0, -1, rd("last_assigned.bin", np.int32)),
                                                                 R["assigned"])),
                "search_by_projection_points": bool(np.array_equal(rd("pts_assigned.bin", np.int32), asg_p_ref))}
    for k in ("search_for_initialization", "search_by_projection_last", "search_by_projection_points"):
        e = g[k]
        e["reference_us_median_1core"] = ref[k]["us_median"]
        e["speedup_vs_reference"] = ref[k]["us_median"] / e["us_median"]
        e["equals_reference"] = same[k] and e["nmatches"] == ref[k]["nmatches"]
        e["compares_per_s"] = e["candidates"] / (e["us_median"] * 1e-6)
        if popc_rate:
            e["frac_of_popc_peak"] = e["compares_per_s"] * 8 / (popc_rate * 1e9)
    out.update(g)
    # ---- class boundary: the SAME driver and timer as the reference leg above, but the ORB_SLAM2::ORBmatcher methods are the
    # product's C++ shim (shim/ORBmatcher_gpu.cc over liborbgpu, compiled against the reference's Frame.h / MapPoint.h): the timed
    # region holds what Tracking waits for -- the gathers under the reference's getters (GetWorldPos, GetDescriptor, Observations),
    # the per-thread device-frame cache lookup, the C-ABI call and the write-back into Frame::mvpMapPoints
    try:
        if os.path.exists(orbref.REF_SHIM_PATH):
            sb = orbref.ShimBackend()
            sb.L.refm_last_call_us.restype = C.c_double
            mks = lambda k, d: Frame(k, d, scale, bounds, backend=sb)
            s1, s2, sc, sl = mks(ka, da), mks(kb, db), mks(ka, da), mks(ka, da)
            node = lambda k, dx, dy: ((k["y"] - dy) // 40).astype(np.int64) * 19 + ((k["x"] - dx) // 40).astype(np.int64)
            na_, nb_ = node(ka, 0, 0), node(kb, 7, 4)          # stand-in for two DBoW2 FeatureVectors: node = 40-px cell
            oa_, ob_ = np.argsort(na_, kind="stable"), np.argsort(nb_, kind="stable")
            ko, ki, fo_, fi = [0], [], [0], []
            for c in np.intersect1d(na_, nb_):
                ki += oa_[na_[oa_] == c].tolist(); fi += ob_[nb_[ob_] == c].tolist()
                ko.append(len(ki)); fo_.append(len(fi))
            kv = np.ones(n, np.uint8)
            calls = {
                "search_for_initialization": lambda B_, f1, f2: ORBmatcher(0.9, True, backend=B_).SearchForInitialization(f1, f2, prev0.copy(), 100),
                "search_by_projection_last": lambda B_, f1, f2: (lambda r: (r["nmatches"], r["assigned"]))(orbref.refm_search_by_projection_last(
                    f1, f2, xyz, has, obs, claimed_last, Tc, Tl, FX3, FY3, CX3, CY3, 40.0, 0.11, 15.0, True, 0.9, True)),
                "search_by_projection_points": lambda B_, f1, f2: ORBmatcher(0.8, backend=B_).SearchByProjectionPoints(
                    f2, in_view, px, py, pxr, lvl, vc, da, 3.0, claimed_pts, obs),
                "search_by_bow": lambda B_, f1, f2: orbref.refm_search_by_bow(f1, f2, ko, ki, fo_, fi, kv, None, False, 0.7, True),
            }
            cls = {}
            for name, call in calls.items():
                tr, ts = [], []
                for _ in range(12):
                    want = call(rb, r1 if name != "search_by_projection_last" else rc, r2 if name != "search_by_projection_last" else rl)
                    tr.append(rb.L.refm_last_call_us())
                for _ in range(reps + 3):
                    got = call(sb, s1 if name != "search_by_projection_last" else sc, s2 if name != "search_by_projection_last" else sl)
                    ts.append(sb.L.refm_last_call_us())
                ts = ts[3:]                                    # the first calls build the thread's device-frame cache
                cls[name] = {"us_median": float(np.median(ts)), "us_min": float(np.min(ts)), "reference_us_median_1core": float(np.median(tr)),
                             "speedup_vs_reference": float(np.median(tr) / np.median(ts)), "nmatches": int(got[0]),
                             "equals_reference": bool(got[0] == want[0] and np.array_equal(got[1], want[1]))}
            cls["how"] = ("ORB_SLAM2::ORBmatcher method calls on Frame / MapPoint objects of the reference's own headers, steady_clock around the "
                          "method: product = shim/ORBmatcher_gpu.cc over liborbgpu (oracle/_ref/liborbmatcher_shim.so), reference = the "
                          "reference's ORBmatcher.cc (liborbmatcher_ref.so), same driver, same inputs, one host thread")
            out["class_boundary"] = cls
    except Exception as exc:   # a checker-side leg: it must not cost the line
        out["class_boundary"] = {"unavailable": repr(exc)}
    out["note"] = ("candidates = (query, feature) pairs whose 256-bit distance is computed (nnz of SURVEY 8d); these calls are latency-bound "
                   "(two kernels + two copies), the POPC fraction is reported for completeness")
    return out


def bench_bow(args, torch, dev, local):
    """DBoW2 transform (Frame::ComputeBoW) on an ORBvoc-shaped tree (k=10, L=6, 1.1 M nodes, levelsup 4):
    `frames` frames x 2000 descriptors per call, device-resident, CUDA events on the vocabulary's stream; the oracle
    (CPU restatement pinned to the reference's DBoW2) timed on a bounded sample beside it."""
    from ar_orbslam2_b200.synth import synth_vocabulary_fast
    from ar_orbslam2_b200.vocabulary import ORBVocabulary
    voc = synth_vocabulary_fast(10, 6, 0)
    G = ORBVocabulary.from_arrays(10, 6, voc["parent"], voc["descriptors"], voc["weight"], voc["word_id"], device=local)
    B, S, steps = 256, 2000, 10
    rng = np.random.default_rng(5)
    leaves = np.nonzero(voc["is_leaf"])[0]
    host = voc["descriptors"][rng.choice(leaves, B * S)] ^ np.packbits(rng.random((B * S, 256), dtype=np.float32) < 0.1, axis=1)
    host = np.ascontiguousarray(host.reshape(B, S, 32))
    d_desc = torch.from_numpy(host).to(dev)
    d_n = torch.full((B,), S, dtype=torch.int32, device=dev)
    bw, fn, ff = (torch.zeros((B, S), dtype=torch.int32, device=dev) for _ in range(3))
    bv = torch.zeros((B, S), dtype=torch.float64, device=dev)
    fo = torch.zeros((B, S + 1), dtype=torch.int32, device=dev)
    nb, nf = torch.zeros(B, dtype=torch.int32, device=dev), torch.zeros(B, dtype=torch.int32, device=dev)
    st = torch.cuda.ExternalStream(G.stream, device=dev)
    torch.cuda.synchronize()
    run = lambda: G.transform_device(d_desc.data_ptr(), d_n.data_ptr(), B, S, 4, S, bw.data_ptr(), bv.data_ptr(), nb.data_ptr(),
                                     fn.data_ptr(), fo.data_ptr(), ff.data_ptr(), nf.data_ptr())
    for _ in range(3):
        run()
    G.sync()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = G.launch_count
    s.record(st)
    for _ in range(steps):
        run()
    e.record(st)
    e.synchronize()
    ms = s.elapsed_time(e) / steps
    r = {"frames_per_s": B / (ms * 1e-3), "features_per_s": B * S / (ms * 1e-3), "ms_per_call": ms,
         "workload": f"{B} frames x {S} descriptors per call, vocabulary k=10 L=6 ({voc['parent'].shape[0]} nodes), levelsup 4",
         "launches": int(G.launch_count - l0), "words_per_frame": float(nb.float().mean().item()),
         "hamming_per_feature": 60}
    # single-frame latency through the blocking host API (what Frame::ComputeBoW costs a tracking thread)
    import time
    one = host[0]
    for _ in range(3):
        G.transform(one, 4)
    t = []
    for _ in range(30):
        t0 = time.perf_counter(); G.transform(one, 4); t.append(time.perf_counter() - t0)
    r["host_call_ms_median"] = float(np.median(t) * 1e3)
    if not args.skip_cpu:
        from oracle import orbref
        O = orbref.RefVocabulary(arrays=voc)
        t0 = time.perf_counter()
        nfr = 8
        for b in range(nfr):
            O.transform(host[b], 4)
        dt = time.perf_counter() - t0
        r["cpu_baseline"] = {"frames_per_s": nfr / dt, "cores": 1, "kind": "port", "sample": f"{nfr} frames x {S} descriptors, one thread"}
        got = G.transform(host[0], 4)
        want = O.transform(host[0], 4)
        r["matches_oracle"] = bool(all(a.tobytes() == b.tobytes() for a, b in zip(got, want)))
    G.close()
    return r


def bench_marker(args):
    """SURVEY 8f-4, the marker pipeline's front end (ref: Marker::setTargetImage / Match, ORB_SLAM2/src/Marker.cc:83-84,107-130):
    OpenCV's cv::ORB (Harris score, 500 features) on the target image and on a camera frame, then the brute-force ratio matcher
    (Marker::searchMatches).  Blocking host calls (pageable image in, key points + descriptors out), wall clock; beside it the
    oracle restatement on one core and -- where cv2 is importable -- cv2's own ORB_create().detectAndCompute, the library the
    reference links (other version), with its threads."""
    import time
    from ar_orbslam2_b200 import cvorb, matcher
    from ar_orbslam2_b200.synth import synth_frame
    r = {"workload": "cv::ORB defaults (500 features, 8 levels, Harris score, FAST 20) on 640x480 and 1920x1080 frames + ratio matcher"}
    for name, (w, h) in (("vga", (640, 480)), ("hd1080", (1920, 1080))):
        orb = cvorb.ORB(max_width=w, max_height=h)
        target, frame = synth_frame(w, h, 100), synth_frame(w, h, 101)
        k1, d1 = orb(target)
        l0 = orb.launch_count
        k2, d2 = orb(frame)
        launches = orb.launch_count - l0
        for _ in range(5):
            orb(frame)
        t = []
        for _ in range(40):
            t0 = time.perf_counter(); orb(frame); t.append(time.perf_counter() - t0)
        leg = {"gpu_ms_median": float(np.median(t) * 1e3), "gpu_ms_min": float(np.min(t) * 1e3), "keypoints": int(len(k2)), "launches": int(launches)}
        t = []
        for _ in range(20):
            t0 = time.perf_counter(); matcher.searchMatches(d1, d2); t.append(time.perf_counter() - t0)
        leg["ratio_match_ms_median"] = float(np.median(t) * 1e3)
        if not args.skip_cpu:
            from oracle import orbref
            t = []
            for _ in range(3 if w > 1000 else 8):
                t0 = time.perf_counter(); rk, rd = orbref.cv_orb(frame); t.append(time.perf_counter() - t0)
            leg["cpu_port_ms_median_1core"] = float(np.median(t) * 1e3)
            leg["equals_oracle"] = bool(rk.tobytes() == k2.tobytes() and np.array_equal(rd, d2))
            try:
                import cv2
                o = cv2.ORB_create()
                o.detectAndCompute(frame, None)
                t = []
                for _ in range(10):
                    t0 = time.perf_counter(); ck, cd = o.detectAndCompute(frame, None); t.append(time.perf_counter() - t0)
                leg["cv2_ms_median"] = float(np.median(t) * 1e3)
                leg["cv2_threads"] = int(cv2.getNumThreads())
                leg["cv2_version"] = cv2.__version__
                leg["equals_cv2"] = bool(len(ck) == len(k2) and np.array_equal(cd, d2) and
                                         all(abs(a.pt[0] - b["x"]) == 0 and abs(a.pt[1] - b["y"]) == 0 and a.response == b["response"] and
                                             a.angle == b["angle"] and a.octave == b["octave"] for a, b in zip(ck, k2)))
            except ImportError:
                leg["cv2_ms_median"] = None
        orb.close()
        r[name] = leg
    return r


def bench_presteps(args, torch, dev):
    """SURVEY 8f-3 pre-steps on device-resident batches of KITTI-shaped frames, CUDA events on the launching stream, against the
    measured HBM peak: cvtColor RGB->grey (reads 3P, writes P) and the stereo-rectification remap (8 map bytes + ~1 source byte
    in, 1 byte out per pixel; the maps are shared by the batch and L2 resident, so its algorithmic bytes are 2P per frame)."""
    from ar_orbslam2_b200.color import cvtColorGray_device
    from ar_orbslam2_b200 import _lib
    import ctypes as C
    W, H, B, steps = 1241, 376, 288, 10
    pitch3, pitch1 = ((3 * W + 15) // 16) * 16, ((W + 15) // 16) * 16
    rgb = torch.randint(0, 256, (B, H, pitch3), dtype=torch.uint8, device=dev)
    grey = torch.zeros((B, H, pitch1), dtype=torch.uint8, device=dev)
    rect = torch.zeros((B, H, pitch1), dtype=torch.uint8, device=dev)
    yy, xx = torch.meshgrid(torch.arange(H, device=dev, dtype=torch.float32), torch.arange(W, device=dev, dtype=torch.float32), indexing="ij")
    mp = ((W + 3) // 4) * 4
    mx = torch.zeros((H, mp), dtype=torch.float32, device=dev); my = torch.zeros((H, mp), dtype=torch.float32, device=dev)
    mx[:, :W] = xx + 2.3 * torch.sin(yy / 40.0) + 0.37
    my[:, :W] = yy + 1.7 * torch.cos(xx / 55.0) + 0.21
    torch.cuda.synchronize()
    st = torch.cuda.Stream(dev)   # an explicit stream: NULL now means the calling thread's own call stream (include/orbgpu.h "Streams")
    L = _lib.lib()
    vp = C.c_void_p
    def cvt():
        cvtColorGray_device(rgb.data_ptr(), W, H, pitch3, H * pitch3, 3, True, B, grey.data_ptr(), pitch1, H * pitch1, stream=st.cuda_stream)
    def rmp():
        _lib.check(L.orb_remap_linear(vp(grey.data_ptr()), W, H, pitch1, C.c_size_t(H * pitch1), vp(mx.data_ptr()), vp(my.data_ptr()), mp,
                                      vp(rect.data_ptr()), W, H, pitch1, C.c_size_t(H * pitch1), B, 1, vp(st.cuda_stream)), "orb_remap_linear")
    out = {}
    hbm_peak, _ = load_peaks()
    for name, fn, bytes_per_frame in (("cvt_color_gray", cvt, 4 * W * H), ("remap_linear", rmp, 2 * W * H)):
        for _ in range(3):
            fn()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(st)
        for _ in range(steps):
            fn()
        e.record(st)
        e.synchronize()
        ms = s.elapsed_time(e) / steps
        gbs = bytes_per_frame * B / (ms * 1e-3) / 1e9
        out[name] = {"frames_per_s": B / (ms * 1e-3), "ms_per_call": ms, "algorithmic_GBps": gbs, "frac_of_hbm_peak": gbs / hbm_peak,
                     "workload": f"{B} frames {W}x{H} per call, device-resident (input {B * H * pitch3 / 1e6:.0f} MB > L2 for cvtColor)"}
    return out


_KEEP = []   # library handles must outlive every pinned buffer that was copied on their streams: torch's pinned-memory
             # allocator records an event on those streams when such a buffer is freed.  They are closed at interpreter exit.


def bench_latency(args):
    """Single-frame latency through the blocking host API (what a tracking thread sees): orb_extract on one 640x480 frame
    (1000 features) and one KITTI-shaped stereo frame (two handles + orb_stereo_match), host image in, host results out."""
    from ar_orbslam2_b200.extractor import ORBextractor
    from ar_orbslam2_b200.stereo import ComputeStereoMatches
    from ar_orbslam2_b200.synth import synth_frame, synth_stereo_pair
    out = {}
    ex = ORBextractor(1000, 1.2, 8, 20, 7, max_width=640, max_height=480)
    img = synth_frame(640, 480, 5)
    for _ in range(5):
        ex(img)
    t = []
    for _ in range(30):
        t0 = time.perf_counter(); ex(img); t.append(time.perf_counter() - t0)
    out["c1_extract_ms_median"] = 1e3 * float(np.median(t))
    out["c1_graph_replays"] = int(ex.graph_replays)   # 0 with ORBGPU_NO_GRAPH=1: the same call as 26 plain launches
    exl = ORBextractor(NFEAT, 1.2, 8, 20, 7, max_width=W, max_height=H)
    exr = ORBextractor(NFEAT, 1.2, 8, 20, 7, max_width=W, max_height=H)
    l, r = synth_stereo_pair(W, H, 9)
    mb = float(np.float32(MBF) / np.float32(FX))
    t = []
    for i in range(35):
        t0 = time.perf_counter()
        kl, dl = exl(l); kr, dr = exr(r)
        ComputeStereoMatches(exl, exr, kl, dl, kr, dr, MBF, mb)
        if i >= 5:
            t.append(time.perf_counter() - t0)
    out["c2_stereo_frame_ms_median"] = 1e3 * float(np.median(t))
    from ar_orbslam2_b200.stereo import StereoFrame
    ex2 = ORBextractor(NFEAT, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=2)
    t = []
    for i in range(35):
        t0 = time.perf_counter()
        StereoFrame(ex2, l, r, MBF, mb)
        if i >= 5:
            t.append(time.perf_counter() - t0)
    out["c2_stereo_frame_single_call_ms_median"] = 1e3 * float(np.median(t))   # orb_stereo_frame: one handle, one synchronisation
    # the same calls from C++ through the drop-in shim classes (what the reference's Tracking thread would wait for)
    exe = os.path.join(ROOT, "ar_orbslam2_b200", "shim", "shim_selftest")
    if os.path.exists(exe):
        import re
        import tempfile
        with tempfile.TemporaryDirectory() as td:
            l.tofile(os.path.join(td, "l.raw")); r.tofile(os.path.join(td, "r.raw"))
            p = subprocess.run([exe, str(W), str(H), os.path.join(td, "l.raw"), os.path.join(td, "r.raw"), str(NFEAT), os.path.join(td, "o"), "time"],
                               capture_output=True, text=True)
        m = re.search(r"latency_us extract=([\d.]+) stereo_three_calls=([\d.]+) stereo_single_call=([\d.]+)", p.stdout)
        if m:
            out["cpp_shim_c2_ms_median"] = {"ORBextractor::operator()": float(m.group(1)) / 1e3,
                                            "two extractions + ORBGPU_ComputeStereoMatches": float(m.group(2)) / 1e3,
                                            "ORBGPU_ExtractStereo": float(m.group(3)) / 1e3}
    out["note"] = "wall clock around blocking calls incl. pageable host copies and the python wrapper; sequential left/right"
    return out


class Lane:
    """One extractor handle with its device-resident outputs and pinned host result buffers (one CUDA stream)."""

    def __init__(self, torch, dev, local, B, host_frames):
        from ar_orbslam2_b200.extractor import ORBextractor
        NF = 2 * B
        self.B, self.NF, self.host = B, NF, host_frames
        self.ex = ORBextractor(NFEAT, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=NF, device=local)
        _KEEP.append(self.ex)
        cap = self.cap = self.ex.default_capacity
        self.stream = torch.cuda.ExternalStream(self.ex.stream, device=dev)
        z = lambda shape, dt: torch.zeros(shape, dtype=dt, device=dev)
        zp = lambda shape, dt: torch.zeros(shape, dtype=dt).pin_memory()
        self.d_in = torch.empty((NF, H, W), dtype=torch.uint8, device=dev)
        self.d = dict(kps=z((NF, cap, 28), torch.uint8), desc=z((NF, cap, 32), torch.uint8), n=z(NF, torch.int32),
                      ur=z((B, cap), torch.float32), dp=z((B, cap), torch.float32), nm=z(B, torch.int32))
        self.h = {k: zp(tuple(v.shape), v.dtype) for k, v in self.d.items()}
        self.mb = float(np.float32(MBF) / np.float32(FX))
        self.torch = torch

    def step_device(self, src, d=None):
        """left+right extraction of B pairs (frames 0..B-1 left, B..2B-1 right) + ComputeStereoMatches, all on the device."""
        from ar_orbslam2_b200.stereo import stereo_match_batch_device
        d, B, cap = d or self.d, self.B, self.cap
        self.ex.extract_batch_device(src.data_ptr(), self.NF, W, H, W, W * H, d["kps"].data_ptr(), d["desc"].data_ptr(), cap,
                                     d["n"].data_ptr())
        stereo_match_batch_device(self.ex, 0, self.ex, B, B, d["kps"].data_ptr(), d["desc"].data_ptr(), d["n"].data_ptr(),
                                  d["kps"][B:].data_ptr(), d["desc"][B:].data_ptr(), d["n"][B:].data_ptr(), cap, MBF, self.mb,
                                  d["ur"].data_ptr(), d["dp"].data_ptr(), d["nm"].data_ptr())

    def enqueue_e2e(self):
        """pinned host frames -> device -> results back in pinned host memory; asynchronous on this lane's stream."""
        with self.torch.cuda.stream(self.stream):
            self.d_in.copy_(self.host, non_blocking=True)
            self.step_device(self.d_in)
            for k in self.d:
                self.h[k].copy_(self.d[k], non_blocking=True)

    def d2h_bytes(self):
        return sum(t.numel() * t.element_size() for t in self.h.values())


class Pipe:
    """End-to-end pipeline around ONE extractor handle: two copy streams feed it.  Frames of step k+1 travel host->device
    and results of step k-1 travel device->host while step k computes (inputs and outputs double-buffered); the compute
    stream never shares the SMs with a second handle's kernels.  NB buffer sets: the host may run NB steps ahead, so the
    upload of step k does not wait for the host to have consumed step k-2 (with two sets the compute stream starved for
    ~0.3 ms per step: 81.5 k frames/s; three sets keep it fed)."""
    NB = 3

    def __init__(self, torch, dev, lane, host):
        self.torch, self.lane, self.host = torch, lane, host
        self.sin, self.sout = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        NB = self.NB
        self.d_in = [torch.empty_like(lane.d_in) for _ in range(NB)]
        self.d = [{k: torch.zeros_like(v) for k, v in lane.d.items()} for _ in range(NB)]
        self.h = [{k: torch.zeros(tuple(v.shape), dtype=v.dtype).pin_memory() for k, v in lane.d.items()} for _ in range(NB)]
        ev = lambda: [torch.cuda.Event() for _ in range(NB)]
        self.in_ready, self.in_free, self.out_ready, self.out_free = ev(), ev(), ev(), ev()
        self.used = [False] * NB

    def enqueue(self, k):
        t, s, cs = self.torch, k % self.NB, self.lane.stream
        if self.used[s]:
            self.sin.wait_event(self.in_free[s])          # the kernels of step k-2 have read this input buffer
        with t.cuda.stream(self.sin):
            self.d_in[s].copy_(self.host, non_blocking=True)
            self.in_ready[s].record(self.sin)
        cs.wait_event(self.in_ready[s])
        if self.used[s]:
            cs.wait_event(self.out_free[s])               # the results of step k-2 have left this output buffer
        self.lane.step_device(self.d_in[s], self.d[s])
        self.in_free[s].record(cs)
        self.out_ready[s].record(cs)
        self.sout.wait_event(self.out_ready[s])
        with t.cuda.stream(self.sout):
            for key in self.d[s]:
                self.h[s][key].copy_(self.d[s][key], non_blocking=True)
            self.out_free[s].record(self.sout)
        self.used[s] = True

    def consume(self, k):
        """Host side: wait until the results of step k are in pinned memory."""
        self.out_free[k % self.NB].synchronize()


def pinned_frames(torch, shape):
    """Pinned host buffer for the input frames.  ORB_BENCH_WC=1: write-combined (cudaHostAllocWriteCombined) -- the CPU only ever
    writes the frames (a camera driver's DMA target), the GPU's PCIe reads then skip the host's cache snoops."""
    if os.environ.get("ORB_BENCH_WC") != "1":
        return torch.empty(shape, dtype=torch.uint8).pin_memory(), "pinned"
    import ctypes
    rt = None
    for name in ("libcudart.so.12", "libcudart.so"):
        try:
            rt = ctypes.CDLL(name)
            break
        except OSError:
            pass
    if rt is None:
        return torch.empty(shape, dtype=torch.uint8).pin_memory(), "pinned (no libcudart for write-combined)"
    n = int(np.prod(shape))
    ptr = ctypes.c_void_p()
    if rt.cudaHostAlloc(ctypes.byref(ptr), ctypes.c_size_t(n), ctypes.c_uint(4)) != 0:
        return torch.empty(shape, dtype=torch.uint8).pin_memory(), "pinned (cudaHostAlloc failed)"
    arr = np.ctypeslib.as_array((ctypes.c_uint8 * n).from_address(ptr.value)).reshape(shape)
    return torch.from_numpy(arr), "pinned write-combined"


def bind_near_gpu(torch, local):
    """Pin this process to the cores next to its GPU (sysfs local_cpulist of the GPU's PCI function) BEFORE the pinned host
    buffers are allocated, so first touch places them on the GPU's NUMA node.  Best effort; returns what was done."""
    try:
        p = torch.cuda.get_device_properties(local)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        path = f"/sys/bus/pci/devices/{bdf}/local_cpulist"
        cpus = set()
        for part in open(path).read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        allowed = os.sched_getaffinity(0)
        use = (cpus & allowed) or allowed
        os.sched_setaffinity(0, use)
        node = open(f"/sys/bus/pci/devices/{bdf}/numa_node").read().strip()
        return {"pci": bdf, "numa_node": int(node), "cpus_bound": len(use), "cpus_allowed": len(allowed)}, allowed
    except Exception as exc:
        return {"unavailable": repr(exc)[:120]}, None


def run_ours(args):
    import torch
    import torch.distributed as dist

    from ar_orbslam2_b200 import _lib

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; liborbgpu has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    numa, affinity0 = bind_near_gpu(torch, local)
    B = args.batch                       # stereo pairs per step (per GPU: weak scaling)
    NF = 2 * B
    pairs = make_pairs(args.distinct, 17 + 1000 * rank)
    host, host_kind = pinned_frames(torch, (NF, H, W))
    hv = host.numpy()
    for b in range(B):
        l, r = pairs[b % len(pairs)]
        hv[b] = l
        hv[B + b] = r
    lanes = [Lane(torch, dev, local, B, host) for _ in range(max(2, args.lanes))]   # lanes: copies overlap compute in the e2e arm
    A = lanes[0]
    ex, stream = A.ex, A.stream
    d_img = host.to(dev)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_max(ms):
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    # ---- device-resident arm: inputs already in HBM, CUDA events on the library's stream
    for _ in range(max(args.warmup, 3)):
        A.step_device(d_img)
    ex.check()
    n_kp = int(A.d["n"].sum().item())
    n_matched = int(A.d["nm"].sum().item())
    L = _lib.lib()
    launches0 = ex.launch_count
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    s.record(stream)
    for _ in range(args.steps):
        A.step_device(d_img)
    e.record(stream)
    e.synchronize()
    barrier()
    ms = reduce_max(s.elapsed_time(e))
    clocks = sampler.stop() if rank == 0 else None
    launches = ex.launch_count - launches0
    ex.check()
    value = NF * args.steps * world / (ms * 1e-3)
    # the same step for at least one second (more steps, not bigger ones): clocks settle, the sampler sees hundreds of samples
    sustained = None
    if ms < 1000.0:
        n_sus = int(math.ceil(1100.0 / (ms / args.steps)))
        sampler2 = ClockSampler(local)
        if rank == 0:
            sampler2.start()
        barrier()
        s.record(stream)
        for _ in range(n_sus):
            A.step_device(d_img)
        e.record(stream)
        e.synchronize()
        barrier()
        ms_sus = reduce_max(s.elapsed_time(e))
        sustained = {"steps": n_sus, "timed_region_s": ms_sus * 1e-3, "value": NF * n_sus * world / (ms_sus * 1e-3), "unit": "frames/s",
                     "ms_per_step": ms_sus / n_sus, "clocks": sampler2.stop() if rank == 0 else None}
    # per-stage device times: a separate, untimed pass with the library's stage events on.  Profiling serialises the
    # schedule (the timed pass above overlaps each level's FAST and blur with the rest of the resize cascade on side
    # streams), so the stage times add up to more than ms_per_step.
    L.orb_extractor_set_profiling(ex._h, 1)
    sp, ep = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sp.record(stream)
    for _ in range(args.steps):
        A.step_device(d_img)
    ep.record(stream)
    ep.synchronize()
    ms_serial = sp.elapsed_time(ep)
    stage_ms = (C.c_double * 6)()
    calls = C.c_int64()
    L.orb_extractor_stage_times(ex._h, stage_ms, C.byref(calls))
    L.orb_extractor_set_profiling(ex._h, 0)
    ex.check()

    # ---- end-to-end arm: every step copies its frames from pinned host memory and its results back to pinned host memory.
    # "pipe" (default): one handle computes, two copy streams move step k+1 in and step k-1 out meanwhile.
    # "lanes": --lanes handles take turns, each doing copy-in, kernels, copy-out on its own stream.
    # Either way a buffer set is synchronised (its host results are consumed) before it is reused.
    e2e_steps = max(6, args.steps)
    s2, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if args.e2e == "pipe":
        pipe = Pipe(torch, dev, A, host)
        NB = pipe.NB
        for k in range(2 * NB):
            if k >= NB:
                pipe.consume(k - NB)
            pipe.enqueue(k)
        barrier()
        t0 = time.perf_counter()
        s2.record(pipe.sin)
        ss, se = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for k in range(e2e_steps):
            if k >= NB:
                pipe.consume(k - NB)
            pipe.enqueue(k)
            if k == 1:
                ss.record(A.stream)      # kernels of step 1 done ...
        se.record(A.stream)              # ... to kernels of the last step done: the pipeline's steady state, fill and drain excluded
        e2.record(pipe.sout)
        e2.synchronize()
        wall_e2e = time.perf_counter() - t0
        barrier()
        results = pipe.h
        pipeline = f"one extractor handle; frames of the next steps in and results of the previous steps out on two copy streams while step k computes ({NB} buffer sets)"
    else:
        NL = len(lanes)
        for i in range(2 * NL):
            lanes[i % NL].stream.synchronize()
            lanes[i % NL].enqueue_e2e()
        barrier()
        t0 = time.perf_counter()
        s2.record(lanes[0].stream)
        for ln in lanes[1:]:
            ln.stream.wait_event(s2)
        for i in range(e2e_steps):
            ln = lanes[i % NL]
            ln.stream.synchronize()
            ln.enqueue_e2e()
        for ln in lanes[1:]:
            join = torch.cuda.Event()
            join.record(ln.stream)
            lanes[0].stream.wait_event(join)
        e2.record(lanes[0].stream)
        e2.synchronize()
        wall_e2e = time.perf_counter() - t0
        barrier()
        results = [ln.h for ln in lanes]
        pipeline = f"{NL} extractor handles take turns; a handle is synchronised before reuse"
    ms_e2e = reduce_max(s2.elapsed_time(e2))
    # copy-only leg: the end-to-end arm's transfers (frames in, results out, both directions at once) with no kernel in between:
    # the rate the host link gives this rank while every other rank does the same -- the ceiling of a copy-bound e2e step
    c_in, c_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    sc, ec, jc = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event()
    h_res = results[0]
    for _ in range(2):
        A.d_in.copy_(host, non_blocking=True)
    barrier()
    sc.record(c_in)
    c_out.wait_event(sc)
    for _ in range(e2e_steps):
        with torch.cuda.stream(c_in):
            A.d_in.copy_(host, non_blocking=True)
        with torch.cuda.stream(c_out):
            for key in A.d:
                h_res[key].copy_(A.d[key], non_blocking=True)
    jc.record(c_out)
    c_in.wait_event(jc)
    ec.record(c_in)
    ec.synchronize()
    barrier()
    ms_copy = reduce_max(sc.elapsed_time(ec))
    steady = None
    if args.e2e == "pipe" and e2e_steps > 2:
        steady = NF * world / (reduce_max(ss.elapsed_time(se)) / (e2e_steps - 2) * 1e-3)
    e2e_value = NF * e2e_steps * world / (ms_e2e * 1e-3)
    h2d, d2h = host.numel(), A.d2h_bytes()
    for r in results:
        assert int(r["n"].sum().item()) == n_kp, "e2e arm produced different keypoint counts"
    # the handles stay open until interpreter exit (ar_orbslam2_b200._lib closes them from atexit): torch's pinned-memory
    # allocator records an event on every stream a pinned buffer was copied on when that buffer is freed, so the
    # library streams must outlive the buffers.

    # the timed step's outputs by hash: every frame and every pair of the e2e arm's last results against the oracle
    gold = golden_sweeps()
    step_hash_ok = None
    if gold and str(rank) in gold["bench"] and args.distinct <= 16:
        from ar_orbslam2_b200.hashing import check_stereo_batch
        step_hash_ok = True
        for r in results:
            bad = check_stereo_batch(gold["bench"][str(rank)][: args.distinct], r["n"].numpy(), r["kps"].numpy(), r["desc"].numpy(),
                                     r["ur"].numpy(), r["dp"].numpy(), r["nm"].numpy())
            if bad:
                raise SystemExit(f"bench.py: rank {rank}: (pair, part) {bad[:6]} of the timed step differ from the oracle's SHA-256")

    if affinity0:
        os.sched_setaffinity(0, affinity0)   # the CPU baselines below use every core again
    extra = {}
    popc = None
    if not os.environ.get("ORB_BENCH_SKIP_HAMMING"):
        extra["c5_hamming"], popc = bench_c5(args, torch, dist, dev, local, rank, world, reduce_max, barrier)
    if not os.environ.get("ORB_BENCH_SKIP_C4"):
        extra["c4_1080p"] = bench_frames_leg(args, torch, dist, dev, local, rank, world, reduce_max, barrier, w=1920, h=1080, nfeat=2000,
                                             total=256, chunk=32, steps=3, golden_key="C4", stages=True,
                                             label="C4: 256 independent frames of 1920x1080, 2000 features, 256/N per GPU")
    if rank == 0 and not os.environ.get("ORB_BENCH_SKIP_C1"):
        extra["c1_mono"] = bench_frames_leg(args, torch, None, dev, local, 0, 1, lambda v: v, torch.cuda.synchronize, w=640, h=480,
                                            nfeat=1000, total=512, chunk=128, steps=5, golden_key="C1",
                                            label="C1: 640x480, 1000 features (mono_tum ORBextractor), 512 frames per step on one GPU")
        extra["latency"] = bench_latency(args)
        extra["c3_tracking"] = bench_c3(args, popc)
        if not os.environ.get("ORB_BENCH_SKIP_BOW"):
            extra["bow_transform"] = bench_bow(args, torch, dev, local)
            extra["presteps"] = bench_presteps(args, torch, dev)
            extra["marker_orb"] = bench_marker(args)

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
        "describe": 2 * S_px + 60 * kp_per_frame,   # blur: read the 8 levels, write the 8 blurred levels; then 60 B per keypoint
    }
    traffic, traffic_file, warp_inst_288 = measured_stage_traffic()
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

    # the ceiling that actually binds (DESIGN.md "Why the HBM fraction is small"): warp instructions of one step, from the named
    # ncu launch list, over the issue slots the step had = 148 SMs x 4 schedulers x SM clock x step time
    issue_frac = None
    if warp_inst_288 and clocks and clocks.get("sm_mhz"):
        issue_frac = warp_inst_288 * NF / 288.0 / (148 * 4 * clocks["sm_mhz"] * 1e6 * (ms / args.steps) * 1e-3)

    def brief(leg, rate_key):   # one config in ~100 bytes, for records that keep `roofline` whole but only the key names of `extra`
        if not leg:
            return None
        o = {"rate": leg.get(rate_key), "e2e": (leg.get("e2e") or {}).get(rate_key),
             "frac": (leg.get("roofline_step") or leg.get("roofline") or {}).get("frac"),
             "cpu": (leg.get("cpu_baseline") or {}).get("value"), "cpu_cores": (leg.get("cpu_baseline") or {}).get("cores")}
        for k in ("all_frames_equal_oracle_sha256", "slice_equals_oracle", "planted_queries_found"):
            if k in leg:
                o[k] = leg[k]
        return o
    c3 = extra.get("c3_tracking") or {}
    configs = {"c1": brief(extra.get("c1_mono"), "frames_per_s"), "c4": brief(extra.get("c4_1080p"), "frames_per_s"),
               "c5": brief(extra.get("c5_hamming"), "gcompares_per_s"),
               "c3_us": {k: (c3.get(k) or {}).get("us_median") for k in ("search_for_initialization", "search_by_projection_last",
                                                                        "search_by_projection_points")} if c3 else None,
               "c3_class_us": {k: {"us": round(v["us_median"], 1), "ref_us": round(v["reference_us_median_1core"], 1), "eq": v["equals_reference"]}
                               for k, v in (c3.get("class_boundary") or {}).items() if isinstance(v, dict) and "us_median" in v} or None,
               "units": "c1/c4 frames/s, c5 Gcompares/s, c3 median microseconds per C-ABI call; c3_class = per ORBmatcher method call "
                        "(product shim vs the reference's ORBmatcher.cc on one core)"}

    line = {
        "metric": "orb_stereo_extract_frames_per_s", "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(B),
        "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e / e2e_steps, "wall_s": wall_e2e, "steps": e2e_steps,
                "pipeline": pipeline,
                "copy_only_value": NF * e2e_steps * world / (ms_copy * 1e-3),
                "host_link_GBps": (h2d + d2h) * e2e_steps / (ms_copy * 1e-3) / 1e9,
                "host_frames_memory": host_kind, "host_link_note": "per rank, all ranks copying at once: this step's H2D + D2H bytes with no kernels, both directions "
                                  "concurrently; e2e cannot exceed min(value, copy_only_value)",
                "numa": numa,
                "steady_state_value": steady,
                "steady_state_note": "same run, compute-stream events after step 1 and after the last step: the timed region above "
                                     "also holds the first upload (no compute to hide behind) and the last download"},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "sustained": sustained,
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                     "frac": achieved / hbm_peak,
                     "traffic": traffic.get(dom, 0) * NF / 288.0 or None,
                     "traffic_source": f"{traffic_file}: ncu dram__bytes_read.sum + dram__bytes_write.sum of the stage's launches in one "
                                       "288-frame step of this command, scaled by frames",
                     "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": alg[dom] * NF, "kernel_ms": kern[dom],
                     "issue_frac": issue_frac,
                     "issue_frac_note": "warp instructions of one step (ncu launch list named in profiles/CURRENT.json) / "
                                        "(148 SMs x 4 schedulers x median SM clock x ms_per_step)",
                     "sustained": sustained, "configs": configs},
        "roofline_stages": {k: {"ms": kern[k], "achieved_GBps": alg[k] * NF / (kern[k] * 1e-3) / 1e9,
                                "frac": alg[k] * NF / (kern[k] * 1e-3) / 1e9 / hbm_peak,
                                "traffic": traffic.get(k, 0) * NF / 288.0 or None} for k in alg},
        "roofline_step": {"bound": "hbm", "achieved": step_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": step_gbs / hbm_peak,
                          "algorithmic_bytes_per_frame": step_alg_bytes, "formula": "P*(1+2S)+60N (SURVEY 8d)"},
        "stage_ms_per_step": dict({STAGES[i]: per_call[i] for i in range(6)},
                                  stereo_match=ms_serial / args.steps - sum(per_call)),
        "ms_per_step_serial_schedule": ms_serial / args.steps,
        "keypoints_per_frame": kp_per_frame, "stereo_matches_per_pair": n_matched / B,
        "timed_step_outputs_equal_oracle_sha256": step_hash_ok,
        "extra": extra,
        "cpu_baseline": {"value": cpu_value, "unit": "frames/s", "cores": cores, "kind": "port",
                         "sample": f"{sample_pairs} synthetic stereo pairs x{reps} through the oracle, one pair per thread"},
    }
    emit_json(line)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400, help="timed steps (the default keeps the timed region above one second; a "
                                                           "shorter run adds a `sustained` block of >= 1.1 s of the same steps)")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=144, help="stereo pairs per step per GPU")
    ap.add_argument("--e2e", default="pipe", choices=["pipe", "lanes"], help="end-to-end arm: copy streams around one handle, or handles taking turns")
    ap.add_argument("--lanes", type=int, default=2, help="e2e arm: steps in flight (each lane = one extractor handle + stream + buffers)")
    ap.add_argument("--distinct", type=int, default=16, help="distinct synthetic pairs tiled into the batch")
    ap.add_argument("--skip-cpu", action="store_true", help="profiling runs: skip the CPU baseline leg")
    ap.add_argument("--db-keyframes", type=int, default=50_000, help="C5: keyframes of 2000 descriptors in the whole database (50 000 = 3.2 GB)")
    args = ap.parse_args()
    # stdout carries exactly ONE line, the JSON: anything libraries print there at the C level (NCCL's version banner under
    # torchrun, for one) goes to stderr instead.  The JSON line is written to the saved descriptor.
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(os.dup(2), "w", buffering=1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
