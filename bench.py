#!/usr/bin/env python
"""bench.py -- headline benchmark of the detection hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" = one pass of the whole path (pyramid -> integrals -> cascade -> groupRectangles) over one
batch of synthetic 1280x720 gray frames with haarcascade_frontalface_alt (the config the metric is
quoted on).  Prints ONE JSON line (rank 0).  `value` = frames/s with frames resident in HBM;
`e2e` = the same through the public C-ABI call with pinned HOST frames (H2D + D2H inside the timed
region; `e2e_pageable` = the same from pageable memory, what a cv::Mat is).  With N > 1 every rank
takes its own frame range and the per-frame lists are gathered to rank 0 inside the library
(ugd_detect_batch_sharded: NCCL send/recv from the device-side result slots).

`parity` in the line is checked IN THIS RUN, on the benchmark frames themselves: the engine's rect
lists (device-resident call and host-frame call) against cv2.detectMultiScale on the frames the
cpu_baseline leg times, and, with N > 1, rank 0's gathered lists against a 1-GPU run of the other
ranks' frames.  A mismatch makes the run fail (exit 3).

`--impl reference` times the reference's own CPU path (cv2 4.13 detectMultiScale, the OpenCV call
detector.cc:25-31 makes) on the host cores, on the same seeded frames and the same `config`.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

W, H = 1280, 720
CASCADE = "alt"
FRAMES_PER_STEP = 256          # 236 MB of input per step (> 126 MB L2); 4 launch groups of 64
GROUP = 64
DISTINCT = 64                  # distinct synthetic frames (seeds 0 .. DISTINCT-1), tiled to FRAMES_PER_STEP; rank r starts ROTATE * r frames in
ROTATE = 5                     # so every rank does the same work per step (weak scaling) on a differently ordered stream
CAP = 64                       # rect slots per frame gathered to rank 0
REF_SAMPLE = 8                 # frames per step of the reference arm (a bounded sample of the step's frames)
METRIC = "720p frames/sec (frontalface_alt)"
WINDOW = {"alt": 20, "alt2": 20, "alt_tree": 20, "default": 24}

# BASELINE.json configs 2-5, run as short legs after the headline so that the driver's records carry them
# (frames per step per GPU, launch group, distinct frames; the 4K pyramid holds 544 MB of integrals per frame)
SUBCONFIGS = [
    {"id": 2, "cascade": "alt2", "size": (1280, 720), "sf": 1.1, "min": 30, "frames": 128, "group": 64, "distinct": 16, "cpu_frames": 8},
    {"id": 3, "cascade": "default", "size": (1920, 1080), "sf": 1.1, "min": 30, "frames": 64, "group": 32, "distinct": 8, "cpu_frames": 4},
    {"id": 4, "cascade": "alt_tree", "size": (1280, 720), "sf": 1.1, "min": 30, "frames": 128, "group": 64, "distinct": 16, "cpu_frames": 8},
    {"id": 5, "cascade": "alt", "size": (3840, 2160), "sf": 1.05, "min": 24, "frames": 16, "group": 8, "distinct": 4, "cpu_frames": 2},
]


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cascade", default=CASCADE)
    ap.add_argument("--frames", type=int, default=FRAMES_PER_STEP)
    ap.add_argument("--size", default=f"{W}x{H}")
    ap.add_argument("--dense-stages", type=int, default=0)
    ap.add_argument("--scale-factor", type=float, default=1.1)
    ap.add_argument("--min-size", type=int, default=30)
    ap.add_argument("--scan-size", default="", help="WxH: BGR frames of --size are decimated (INTER_AREA) to this scan size "
                                                    "and run through detectFaces, as the tracker does with conf/*.cfg")
    ap.add_argument("--group", type=int, default=GROUP, help="frames per launch group")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the short legs for BASELINE configs 2-5")
    return ap.parse_args()


def make_frames(synth, w, h, n, seed0, distinct=DISTINCT):
    base = synth.gray_batch(w, h, min(distinct, n), seed0=seed0)
    reps = (n + len(base) - 1) // len(base)
    return np.ascontiguousarray(np.concatenate([base] * reps)[:n])


def geometry(w, h, win, sf, ms):
    """Pyramid geometry in plain Python (running product in double, float32 scale, float32 division,
    half-even rounding) -- so that the reference arm can print the same `config` without loading the CUDA library."""
    f32 = np.float32
    levels, level_px, int_entries, f = 0, 0, 0, 1.0
    while True:
        if int(np.rint(win * f)) > w or int(np.rint(win * f)) > h:
            break
        s = f32(f)
        ww = int(np.rint(f32(win) * s))
        f *= sf
        if ww > min(w, h):
            break
        if ww < ms:
            continue
        sw, sh = int(np.rint(f32(w) / s)), int(np.rint(f32(h) / s))
        if sw + 1 - win <= 0 or sh + 1 - win <= 0:
            continue
        levels += 1
        level_px += sw * sh
        int_entries += (sw + 1) * (sh + 1)
    return levels, level_px, int_entries


def workload_config(cascade_file, w, h, sf, ms, nF, group, world, win, scan=None, distinct=DISTINCT):
    """The `config` object: identical in both arms (same frames, same parameters, same words)."""
    levels, level_px, int_entries = geometry(scan[0] if scan else w, scan[1] if scan else h, win, sf, ms)
    head = (f"tracker front end: synthetic {w}x{h} BGR frames -> INTER_AREA to {scan[0]}x{scan[1]} -> detectFaces (skin mask), "
            f"{cascade_file}, " if scan else f"{cascade_file} on synthetic {w}x{h} gray frames, ")
    tail = " (BASELINE configs[0]/[1] frame+params)" if (w, h, sf, ms) == (1280, 720, 1.1, 30) and not scan else ""
    return {"workload": head + f"scale {sf}, minNeighbors 2, minSize {ms}x{ms}" + tail,
            "frames_per_step_per_gpu": nF, "launch_group": group, "levels": levels,
            "frames": f"seeds 0..{min(distinct, nF) - 1} (SURVEY A.7 generator), tiled to {nF} per rank per step; rank r starts {ROTATE}*r frames in "
                      f"(same work on every rank, different order)",
            "l2": f"inputs larger than L2: {nF * w * h * (3 if scan else 1) / 1e6:.0f} MB of frames per step, "
                  f"{group * (level_px + 8 * int_entries) / 1e6:.0f} MB of integrals per launch group",
            "parallelism": f"frames sharded over {world} rank(s), lists gathered to rank 0 (NCCL inside the library)"}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return d.get("hbm_gbs", 6650.0), d.get("sm_max_mhz", 1965.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, 1965.0, "fallback (B200_PROFILING.md)"


def traffic_record(kernel, key):
    """DRAM bytes per launch group from a committed ncu capture (profiles/traffic.json), never a constant in this file.
    Returns (bytes or None, source string)."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(p) as f:
            rec = json.load(f).get(kernel, {}).get(key)
    except (OSError, ValueError):
        rec = None
    if not rec:
        return None, "no ncu capture recorded for this workload (profiles/traffic.json)"
    return rec["dram_bytes"], f"{rec['source']} (build {rec['build']})"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.t0 = self.t1 = None                      # wall-clock bounds of the timed region
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--id={gpu_index}", f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush(); self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f.read().splitlines():
            c = [t.strip() for t in line.split(",")]
            if len(c) < 9:
                continue
            try:
                import datetime
                ts = datetime.datetime.strptime(c[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
                if self.t0 is not None and not (self.t0 - 0.05 <= ts <= self.t1 + 0.05):
                    continue                          # sample taken outside the timed region
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        if sm:
            out = {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}
        return out


def rectset(a):
    return sorted(tuple(int(v) for v in r) for r in np.asarray(a).reshape(-1, 4))


def cpu_reference_run(cv2ref, frames, cascade_name, budget_s, max_frames, sf=1.1, ms=(30, 30)):
    """The reference's CPU path: cv2.CascadeClassifier.detectMultiScale(gray, 1.1, 2, SCALE_IMAGE, (30,30)) with all
    host threads, frame i of `frames` in order.  Returns (frames/s, detections per frame timed, seconds, threads)."""
    import cv2
    cv2.setNumThreads(os.cpu_count() or 1)
    cc = cv2ref.load(cv2ref.cascade_path(cascade_name))
    cv2ref.detect_multiscale_cv2(cc, frames[0], sf, 2, ms)   # warm-up
    dets, t0 = [], time.perf_counter()
    while len(dets) < min(max_frames, len(frames)):
        dets.append(cv2ref.detect_multiscale_cv2(cc, frames[len(dets)], sf, 2, ms))
        if time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    return len(dets) / dt, dets, dt, cv2.getNumThreads()


def run_reference(args, rank, world):
    """--impl reference: OpenCV CPU path on the host cores, rank 0 only, on the frames of the repo arm's rank 0."""
    if rank != 0:
        return
    from oracle import cv2_ref
    ge.load_package()                                       # only for the synthetic-frame generator; the .so is not loaded
    synth = __import__("importlib").import_module(ge.PKG_NAME + ".synth")
    import cv2
    w, h = map(int, args.size.split("x"))
    nF = args.frames
    sf, ms = args.scale_factor, (args.min_size, args.min_size)
    frames = make_frames(synth, w, h, min(DISTINCT, nF), seed0=0)          # == rank 0's frames of the repo arm (seeds i, no rotation)
    cv2.setNumThreads(os.cpu_count() or 1)
    path = cv2_ref.cascade_path(args.cascade)
    cc = cv2_ref.load(path)
    win = cc.getOriginalWindowSize()[0]
    for _ in range(args.warmup):
        cv2_ref.detect_multiscale_cv2(cc, frames[0], sf, 2, ms)
    t0 = time.perf_counter()
    k = 0
    for _ in range(args.steps):                              # each step: a bounded sample of the step's frames
        for _ in range(REF_SAMPLE):
            cv2_ref.detect_multiscale_cv2(cc, frames[k % len(frames)], sf, 2, ms)
            k += 1
    dt = time.perf_counter() - t0
    fps = k / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8/int32/f32", "data": "synthetic",
        "config": workload_config(os.path.basename(path), w, h, sf, ms[0], nF, min(args.group, nF), world, win),
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cv2.getNumThreads(), "kind": "reference",
                         "sample": f"{REF_SAMPLE} of the step's {nF} frames per step x {args.steps} steps (rank 0's frames, in order), "
                                   f"cv2 {cv2.__version__} detectMultiScale, {os.cpu_count()} host cpus"},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


class Runner:
    """One workload on this rank's GPU: detector, frames (device, pinned, pageable), timed steps, parity."""

    def __init__(self, env, cascade, w, h, sf, ms, nF, group, distinct, scan=None, dense_stages=0):
        import torch
        self.env, self.torch = env, torch
        ug, synth, cv2_ref = env["ug"], env["synth"], env["cv2_ref"]
        self.synth = synth
        self.cascade, self.w, self.h, self.sf, self.ms, self.nF, self.scan = cascade, w, h, sf, (ms, ms), nF, scan
        self.group = min(group, nF)
        self.distinct = min(distinct, nF)
        self.casc = ug.Cascade(cv2_ref.cascade_path(cascade))
        prm = ug.default_params(dense_stages=dense_stages, scale_factor=sf, min_size=self.ms)
        self.det = ug.Detector(self.casc, prm, device=env["local"], max_size=scan or (w, h), max_batch=self.group)
        self.det.set_stream(env["stream"].cuda_stream)
        if env["world"] > 1:
            env["sharding"].init_comm(self.det, env["rank"], env["world"])
        rank = env["rank"]
        if scan:
            base = np.stack([synth.bgr_frame(w, h, synth.default_nfaces(w, h), i) for i in range(self.distinct)])
        else:
            base = synth.gray_batch(w, h, self.distinct, seed0=0)
        base = np.roll(base, -ROTATE * rank, axis=0)        # rank r's frame j is distinct frame (j + ROTATE * r) mod distinct
        frames_np = np.ascontiguousarray(np.concatenate([base] * ((nF + len(base) - 1) // len(base)))[:nF])
        self.pageable = frames_np                            # plain numpy allocation: pageable, like a cv::Mat
        self.host = torch.from_numpy(frames_np).pin_memory()
        self.dev = self.host.cuda(non_blocking=False)
        self.host_np = self.host.numpy()

    def close(self):
        self.det.close()

    # one step = this rank's nF frames through the path (+ the gather to rank 0 when world > 1)
    def _run(self, ptr, is_device):
        det, env = self.det, self.env
        if self.scan:
            return det.detect_batch_scaled_ptr(ptr, is_device, self.nF, self.w, self.h, self.scan, cap=CAP)
        if env["world"] > 1:
            return det.detect_batch_sharded(ptr, is_device, False, self.nF, self.w, self.h, self.nF, 0, CAP)
        return det._batch(__import__("ctypes").c_void_p(ptr), int(is_device), False, self.nF, self.w, self.h, CAP)

    def step_device(self):
        return self._run(self.dev.data_ptr(), True)

    def step_host(self):
        return self._run(self.host_np.ctypes.data, False)

    def step_pageable(self):
        return self._run(self.pageable.ctypes.data, False)

    def single_frame_latency(self, n=60):
        """One frame per call, the way the reference's tracker calls detectFaces (face_tracker.cc:295): wall-clock time of the
        synchronous ugd_detect_gray / ugd_detect_bgr call on a pageable frame, median of n calls after 10 warm-up calls."""
        out = {}
        w, h = self.scan or (self.w, self.h)
        frames = {"detect_gray_ms": self.synth.gray_frame(w, h, self.synth.default_nfaces(w, h), 7),
                  "detect_bgr_ms": self.synth.bgr_frame(w, h, self.synth.default_nfaces(w, h), 7)}
        for key, f in frames.items():
            fn = self.det.detect_gray if f.ndim == 2 else self.det.detect_bgr
            for _ in range(10):
                fn(f)
            ts = []
            for _ in range(n):
                t0 = time.perf_counter()
                fn(f)
                ts.append((time.perf_counter() - t0) * 1e3)
            out[key] = statistics.median(ts)
        out["frame"] = f"{w}x{h}, pageable host memory, one synchronous call per frame"
        return out

    def timed(self, fn, steps, warmup, sample_clocks=False):
        torch, env = self.torch, self.env
        dist = env["dist"]
        sampler = ClockSampler(env["local"]) if sample_clocks and env["rank"] == 0 else None   # nvidia-smi needs ~1 s to start
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        if env["world"] > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if sampler:
            sampler.t0 = time.time()
        e0.record(env["stream"])
        last = None
        for _ in range(steps):
            last = fn()
        e1.record(env["stream"])
        torch.cuda.synchronize()
        if sampler:
            sampler.t1 = time.time()
        if env["world"] > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        clocks = sampler.stop() if sampler else None
        if env["world"] > 1:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, clocks, last

    def own_lists(self, result):
        """This rank's per-frame lists out of a step's result (rank 0 of a sharded call holds every rank's)."""
        if self.env["world"] > 1 and not self.scan and self.env["rank"] == 0:
            return result[0]
        return result

    def parity(self, results, cpu_frames, budget_s):
        """Engine vs cv2 on THESE frames (rank 0): `results` = {"device": lists, "host": lists, ...} of the timed steps.
        Returns (parity dict, cpu_baseline dict)."""
        env, cv2_ref = self.env, self.env["cv2_ref"]
        import cv2
        sf, ms = self.sf, self.ms
        if self.scan:
            cv2.setNumThreads(os.cpu_count() or 1)
            cc = cv2_ref.load(cv2_ref.cascade_path(self.cascade))
            dets, t0 = [], time.perf_counter()
            while len(dets) < min(cpu_frames, self.distinct) and time.perf_counter() - t0 < budget_s:
                small = cv2.resize(self.host_np[len(dets)], self.scan, interpolation=cv2.INTER_AREA)
                dets.append(cv2_ref.detect_faces_cv2(cc, small, sf))
            cdt = time.perf_counter() - t0
            cfps, cthreads, what = len(dets) / cdt, cv2.getNumThreads(), "cv2 INTER_AREA resize + detectFaces"
        else:
            cfps, dets, cdt, cthreads = cpu_reference_run(cv2_ref, self.host_np[:self.distinct], self.cascade, budget_s, cpu_frames, sf, ms)
            what = "cv2 detectMultiScale"
        mism, checked, total_rects = [], 0, 0
        for kind, res in results.items():
            mine = self.own_lists(res)
            for i, want in enumerate(dets):
                # the step's frames are the distinct ones tiled: frame i and every copy of it must equal cv2's list
                for j in range(i, self.nF, self.distinct):
                    checked += 1
                    if rectset(mine[j]) != rectset(want):
                        mism.append({"call": kind, "frame": j, "engine": rectset(mine[j]), "cv2": rectset(want)})
                total_rects += len(want)
        par = {"frames_checked": checked, "distinct_frames": len(dets), "calls": sorted(results), "mismatches": len(mism),
               "rects_compared": total_rects, "against": f"cv2 {cv2.__version__} {what} on the same frames (sets per frame, exact)"}
        if mism:
            par["first_mismatches"] = mism[:3]
        cpu = {"value": cfps, "unit": "frames/s", "cores": cthreads, "kind": "reference",
               "sample": f"{len(dets)} of the step's {self.nF} frames in {cdt:.1f} s, {what}, {os.cpu_count()} host cpus"}
        return par, cpu

    def gather_parity(self, result, per_rank=4):
        """world > 1, rank 0: the gathered lists of the OTHER ranks against a 1-GPU run of their frames here."""
        env = self.env
        ug, synth = env["ug"], env["synth"]
        single = ug.Detector(self.casc, ug.default_params(scale_factor=self.sf, min_size=self.ms), device=env["local"],
                             max_size=(self.w, self.h), max_batch=per_rank)
        checked, mism = 0, []
        base = synth.gray_batch(self.w, self.h, self.distinct, seed0=0)
        for r in range(1, env["world"]):
            if len(result[r]) != self.nF:
                mism.append({"rank": r, "error": f"{len(result[r])} lists gathered, want {self.nF}"})
                continue
            fr = np.roll(base, -ROTATE * r, axis=0)[:per_rank]
            want = single.detect_batch(np.ascontiguousarray(fr), cap=CAP)
            for i in range(per_rank):
                for j in range(i, self.nF, self.distinct):
                    checked += 1
                    if rectset(result[r][j]) != rectset(want[i]):
                        mism.append({"rank": r, "frame": j, "gathered": rectset(result[r][j]), "single_gpu": rectset(want[i])})
        single.close()
        out = {"gathered_frames_checked": checked, "gathered_mismatches": len(mism)}
        if mism:
            out["first_gather_mismatches"] = mism[:3]
        return out


def launches_per_group(runner, counters):
    """Kernels of ours per launch group: resize+colsum, carry, integral bands, scan (one launch per grid-step class
    present), group init / link / finish; + (area,) preprocess, equalize on the BGR path."""
    n = 3 + 3
    sc, lw, lh, stp = runner.det.plan_info(*(runner.scan or (runner.w, runner.h)))
    n += len(set(int(s) for s in stp))
    if runner.scan:
        sw, sh = runner.scan
        fused = (2 * sw, 2 * sh) == (runner.w, runner.h) and sw % 4 == 0     # the 2x2 decimation runs inside k_preprocess: no k_area launch
        n += 2 if fused else 3
    return n


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the engine has no CPU fallback")
    torch.cuda.set_device(local)
    if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION", "WARN"):
        # keep stdout to the one JSON line: NCCL prints "NCCL version ..." to stdout at VERSION *and* at WARN (an nccl.conf
        # may ask for VERSION); warnings go to stderr instead
        os.environ["NCCL_DEBUG"] = "WARN"
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ug = ge.load_package()
    imp = __import__("importlib").import_module
    from oracle import cv2_ref
    env = {"ug": ug, "synth": imp(ge.PKG_NAME + ".synth"), "sharding": imp(ge.PKG_NAME + ".sharding"), "cv2_ref": cv2_ref,
           "rank": rank, "world": world, "local": local, "dist": dist, "stream": torch.cuda.current_stream()}

    w, h = map(int, args.size.split("x"))
    nF = args.frames
    sf, ms = args.scale_factor, args.min_size
    scan = tuple(map(int, args.scan_size.split("x"))) if args.scan_size else None
    R = Runner(env, args.cascade, w, h, sf, ms, nF, args.group, DISTINCT, scan, args.dense_stages)
    det, GROUP_ = R.det, R.group
    warm = max(args.warmup, 3)

    ms_dev, clocks, r_dev = R.timed(R.step_device, args.steps, warm, sample_clocks=True)
    ms_host, _, r_host = R.timed(R.step_host, args.steps, 3)
    ctr = det.counters()                                   # counters of the last step (nF frames)
    ms_page, _, r_page = R.timed(R.step_pageable, max(2, args.steps // 4), 1)
    n_page = max(2, args.steps // 4)

    # per-kernel device time (CUDA events on the launch stream, inside the library) -> roofline
    det.set_profiling(True)
    ktimes = {k: [0.0, 0] for k in ug.K_NAMES}
    for _ in range(max(2, min(args.steps, 3))):
        R.step_device()
        for k, (t, l) in det.kernel_times().items():
            ktimes[k][0] += t; ktimes[k][1] += l
    det.set_profiling(False)
    latency = R.single_frame_latency() if rank == 0 else None

    line = None
    failed = False
    if rank == 0:
        hbm_peak, sm_max, peak_src = peaks()
        total_frames = nF * world * args.steps
        fps = total_frames / (ms_dev * 1e-3)
        fps_e2e = total_frames / (ms_host * 1e-3)
        fps_page = nF * world * n_page / (ms_page * 1e-3)
        level_px = ctr["level_pixels"]
        pw, ph = scan or (w, h)                                 # size the pyramid is built from
        src_px = pw * ph
        sc, lw, lh, stp = det.plan_info(pw, ph)
        int_entries = int(sum((int(a) + 1) * (int(b) + 1) for a, b in zip(lw, lh)))
        weak_per_frame = ctr["weak_evals"] / max(ctr["frames"], 1)
        win_per_frame = ctr["windows_evaluated"] / max(ctr["frames"], 1)
        # algorithmic bytes per frame (DESIGN.md "Kernels and rooflines")
        alg = {
            "k_resize": src_px + level_px,                         # 1 B/source px read once + 1 B/level px written (+ the strip column sums: noise)
            "k_integral": level_px + 8 * int_entries,              # 1 B read + two int32 planes written
            "k_scan": 0.0,                                         # L1/shared-bound, see roofline below
            "k_area": 3 * (w * h + src_px) if scan else 0.0,       # BGR: 3 B read per source px + 3 B written per scan px
            "k_preprocess": 5 * src_px if scan else 0.0,           # 3 B read + gray + mask written
            "k_equalize": 3 * src_px if scan else 0.0,             # gray + mask read, equalised gray written
        }
        # cascade: 4-byte gathers; merged corner lists need 7.5 per weak classifier on average for alt
        # (6/8/9 distinct corners instead of 8/12), + 8 for the variance norm of every window
        gather_bytes = weak_per_frame * 4 * 7.5 + win_per_frame * 4 * 8
        kern = {}
        for k, (t, l) in ktimes.items():
            if l == 0:
                continue
            kern[k] = {"ms_per_launch": t / l, "launches": l}
        total_k = sum(v["ms_per_launch"] for v in kern.values())
        for k, v in kern.items():
            v["share"] = v["ms_per_launch"] / total_k if total_k else 0.0
            if alg.get(k):
                gbs = alg[k] * GROUP_ / (v["ms_per_launch"] * 1e-3) / 1e9
                v.update({"bound": "hbm", "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak})
        # the image stage as ONE unit: pyramid + both integral planes, whatever kernels it takes (algorithmic bytes:
        # source read once, planes written once; the pyramid itself is an intermediate, not an output of the path)
        img_ms = sum(kern[k]["ms_per_launch"] for k in ("k_resize", "k_carry", "k_integral") if k in kern)
        int_ms = sum(kern[k]["ms_per_launch"] for k in ("k_carry", "k_integral") if k in kern)   # (the strip column sums are computed inside k_resize)
        key = f"{args.cascade} {w}x{h} group {GROUP_}"
        stage = {}
        if img_ms:
            g1 = (src_px + 8 * int_entries) * GROUP_ / (img_ms * 1e-3) / 1e9
            tb1, tsrc1 = traffic_record("image_stage", key)
            stage["image_stage"] = {"traffic": tb1, "traffic_source": tsrc1,"what": "pyramid resize + S/SQ integrals of one launch group, all kernels", "ms": img_ms,
                                    "algorithmic_bytes": (src_px + 8 * int_entries) * GROUP_, "achieved": g1, "peak": hbm_peak,
                                    "unit": "GB/s", "frac": g1 / hbm_peak}
        if int_ms:
            g2 = (level_px + 8 * int_entries) * GROUP_ / (int_ms * 1e-3) / 1e9
            tb, tsrc = traffic_record("integral", key)
            stage["integral_unit"] = {"what": "S/SQ integrals of one launch group: k_carry + k_integral_bands (the strip column sums are "
                                              "produced by the fused resize kernel)", "ms": int_ms,
                                      "algorithmic_bytes": (level_px + 8 * int_entries) * GROUP_, "achieved": g2, "peak": hbm_peak,
                                      "unit": "GB/s", "frac": g2 / hbm_peak, "traffic": tb, "traffic_source": tsrc}
        if scan:
            # the BGR front end of the tracker's workload as ONE unit: INTER_AREA decimation + gray / skin mask / histogram +
            # equalisation (algorithmic bytes: the full-size BGR frame read once; gray, mask and the equalised image written once)
            fe_ms = sum(kern[k]["ms_per_launch"] for k in ("k_area", "k_preprocess", "k_equalize") if k in kern)
            fe_bytes = (3 * w * h + 3 * scan[0] * scan[1]) * GROUP_
            if fe_ms:
                g3 = fe_bytes / (fe_ms * 1e-3) / 1e9
                tbf, tsf = traffic_record("bgr_front_end", f"{args.cascade} {w}x{h} -> {scan[0]}x{scan[1]} group {GROUP_}")
                stage["bgr_front_end"] = {"what": "INTER_AREA decimation + BGR->gray / skin mask / histogram + equalizeHist of one launch group"
                                                  + (" (the 2x2 decimation runs inside k_preprocess)" if "k_area" not in kern else ""),
                                          "ms": fe_ms, "algorithmic_bytes": fe_bytes, "achieved": g3, "peak": hbm_peak, "unit": "GB/s",
                                          "frac": g3 / hbm_peak, "traffic": tbf, "traffic_source": tsf}
        casc_ms = kern["k_scan"]["ms_per_launch"] if "k_scan" in kern else 0.0   # both scan launches of a group
        l1_nominal = 148 * 128 * sm_max * 1e6 / 1e9            # GB/s: 148 SMs x 128 B/clk x max SM clock
        try:
            import ctypes as C
            g_, b_ = C.c_double(), C.c_double()
            rc = ug.lib().ugd_measure_lds_peak(local, C.byref(g_), C.byref(b_))
            lds_meas, lds_bpc = (g_.value, b_.value) if rc == 0 else (None, None)
        except Exception:                                       # noqa: BLE001 -- the probe is evidence, not the product
            lds_meas, lds_bpc = None, None
        l1_peak = lds_meas or l1_nominal
        casc_gbs = gather_bytes * GROUP_ / (casc_ms * 1e-3) / 1e9 if casc_ms else 0.0
        dominant = max(kern, key=lambda k: kern[k]["ms_per_launch"]) if kern else None
        if dominant == "k_scan":
            tb, tsrc = traffic_record("k_scan", key)
            roof = {"kernel": "k_scan (window scan + all cascade stages)", "bound": "l1", "achieved": casc_gbs, "peak": l1_peak,
                    "unit": "GB/s", "frac": casc_gbs / l1_peak,
                    "peak_source": ("measured in this run: conflict-free ld.shared.b32 stream on every SM (ugd_measure_lds_peak, "
                                    f"{lds_bpc:.1f} B/clk/SM at the rated clock)" if lds_meas else "nominal 148 SM x 128 B/clk x max SM clock"),
                    "peak_nominal": l1_nominal, "frac_of_nominal": casc_gbs / l1_nominal,
                    "traffic": tb, "traffic_source": tsrc,
                    "hbm": {"algorithmic_bytes": 8 * int_entries * GROUP_,
                            "achieved": 8 * int_entries * GROUP_ / (casc_ms * 1e-3) / 1e9 if casc_ms else None,
                            "peak": hbm_peak, "unit": "GB/s",
                            "frac": 8 * int_entries * GROUP_ / (casc_ms * 1e-3) / 1e9 / hbm_peak if casc_ms else None},
                    "note": "gather/threshold work: bounded by shared-memory (LSU) throughput, not HBM; algorithmic bytes = 4 B x "
                            "(8 norm corners per window + 7.5 corners per weak classifier); the `hbm` sub-object gives the same "
                            "launch against the HBM roofline (planes read once) to show how far from HBM-bound it is"}
        else:
            v = kern.get(dominant, {})
            roof = {"kernel": dominant, "bound": v.get("bound", "hbm"), "achieved": v.get("achieved"), "peak": hbm_peak,
                    "unit": "GB/s", "frac": v.get("frac"), "traffic": None}
        iu = stage.get("integral_unit", {})
        roof_hbm = {"kernel": "integral S+SQ (all of its launches)", "bound": "hbm", "achieved": iu.get("achieved"),
                    "peak": hbm_peak, "unit": "GB/s", "frac": iu.get("frac"), "traffic": iu.get("traffic"),
                    "traffic_source": iu.get("traffic_source"), "peak_source": peak_src}

        results = {"device": r_dev, "host": r_host, "pageable": r_page}
        parity, cpu = ({"frames_checked": 0, "mismatches": 0, "skipped": "--no-cpu-baseline"}, None)
        if not args.no_cpu_baseline:
            parity, cpu = R.parity(results, 128, 12.0)
        parity["near_threshold"] = ctr["near_threshold"]
        if world > 1 and not scan:
            parity.update(R.gather_parity(r_dev))
        failed = bool(parity.get("mismatches") or parity.get("gathered_mismatches"))
        groups = (nF + GROUP_ - 1) // GROUP_
        lpg = launches_per_group(R, ctr)
        line = {
            "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8/int32/f32", "data": "synthetic",
            "config": workload_config(os.path.basename(cv2_ref.cascade_path(args.cascade)), w, h, sf, ms, nF, GROUP_, world,
                                      R.casc.info.win_w, scan),
            "windows_evaluated_per_s": win_per_frame * fps,
            "weak_classifiers_per_s": weak_per_frame * fps,
            "e2e": {"value": fps_e2e, "unit": "frames/s", "h2d_bytes_per_step": nF * w * h * (3 if scan else 1),
                    "d2h_bytes_per_step": nF * (CAP * 16 + 4 + 64), "ms_per_step": ms_host / args.steps, "host_memory": "pinned"},
            "e2e_pageable": {"value": fps_page, "unit": "frames/s", "ms_per_step": ms_page / n_page, "steps": n_page,
                             "host_memory": "pageable (a plain allocation, what the tracker's cv::Mat is: face_tracker.cc:112-114)"},
            "gpu_launches": (groups * lpg + (1 if world > 1 else 0)) * args.steps,
            "single_frame_latency": latency,
            "parity": parity,
            "roofline": roof, "roofline_hbm": roof_hbm, "stages": stage, "kernels": kern,
            "cpu_baseline": cpu, "clocks": clocks,
            "counters_last_step": ctr,
        }
    R.close()
    del R

    # ---- BASELINE configs 2-5: short legs at this run's N, same rules (3 warm-up steps, device-timed, max over ranks) ----
    if not args.no_configs and not scan and (args.cascade, w, h) == (CASCADE, W, H):
        legs = []
        for c in SUBCONFIGS:
            cw, chh = c["size"]
            Rc = Runner(env, c["cascade"], cw, chh, c["sf"], c["min"], c["frames"], c["group"], c["distinct"])
            steps_c = 3
            ms_d, _, rd = Rc.timed(Rc.step_device, steps_c, 3)
            ms_h, _, rh = Rc.timed(Rc.step_host, steps_c, 1)
            cctr = Rc.det.counters()
            if rank == 0:
                par, cpu_c = ({"frames_checked": 0, "mismatches": 0, "skipped": "--no-cpu-baseline"}, None)
                if not args.no_cpu_baseline:
                    par, cpu_c = Rc.parity({"device": rd, "host": rh}, c["cpu_frames"], 8.0)
                if world > 1:
                    par.update(Rc.gather_parity(rd, per_rank=1))
                failed = failed or bool(par.get("mismatches") or par.get("gathered_mismatches"))
                tot = c["frames"] * world * steps_c
                legs.append({"baseline_config": c["id"],
                             "config": workload_config(os.path.basename(cv2_ref.cascade_path(c["cascade"])), cw, chh, c["sf"], c["min"],
                                                       c["frames"], Rc.group, world, Rc.casc.info.win_w, distinct=c["distinct"]),
                             "value": tot / (ms_d * 1e-3), "unit": "frames/s", "n_gpus": world, "steps": steps_c, "warmup": 3,
                             "ms_per_step": ms_d / steps_c,
                             "e2e": {"value": tot / (ms_h * 1e-3), "unit": "frames/s", "h2d_bytes_per_step": c["frames"] * cw * chh,
                                     "d2h_bytes_per_step": c["frames"] * (CAP * 16 + 4 + 64)},
                             "windows_evaluated_per_frame": cctr["windows_evaluated"] / max(cctr["frames"], 1),
                             "parity": par, "cpu_baseline": cpu_c})
            Rc.close()
            del Rc
        if rank == 0:
            line["configs"] = legs

    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    if failed:
        sys.stderr.write("bench.py: PARITY MISMATCH on the benchmark frames -- see `parity` in the JSON line\n")
        sys.exit(3)


if __name__ == "__main__":
    main()
