#!/usr/bin/env python
"""bench.py -- headline benchmark of the detection hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" = one pass of the whole path (pyramid -> integrals -> cascade -> groupRectangles) over one
batch of synthetic 1280x720 gray frames with haarcascade_frontalface_alt (the config the metric is
quoted on).  Prints ONE JSON line (rank 0).  `value` = frames/s with frames resident in HBM;
`e2e` = the same through the public C-ABI call with pinned HOST frames (H2D + D2H inside the timed
region).  `--impl reference` times the reference's own CPU path (cv2 4.13 detectMultiScale, the
OpenCV call detector.cc:25-31 makes) on the host cores.
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
DISTINCT = 32                  # distinct synthetic frames, tiled to FRAMES_PER_STEP
CAP = 64                       # rect slots per frame gathered to rank 0
METRIC = "720p frames/sec (frontalface_alt)"


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
    return ap.parse_args()


def make_frames(synth, w, h, n, seed0):
    base = synth.gray_batch(w, h, min(DISTINCT, n), seed0=seed0)
    reps = (n + len(base) - 1) // len(base)
    return np.ascontiguousarray(np.concatenate([base] * reps)[:n])


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return d.get("hbm_gbs", 6650.0), d.get("sm_max_mhz", 1965.0), "measured"
    return 6650.0, 1965.0, "fallback"


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


def cpu_reference_run(cv2ref, frames, cascade_name, budget_s, max_frames, sf=1.1, ms=(30, 30)):
    """The reference's CPU path: cv2.CascadeClassifier.detectMultiScale(gray, 1.1, 2, SCALE_IMAGE, (30,30))
    with all host threads.  Returns (frames/s, frames timed, seconds, threads)."""
    import cv2
    cv2.setNumThreads(os.cpu_count() or 1)
    cc = cv2ref.load(cv2ref.cascade_path(cascade_name))
    cv2ref.detect_multiscale_cv2(cc, frames[0], sf, 2, ms)   # warm-up
    n, t0 = 0, time.perf_counter()
    while n < max_frames:
        cv2ref.detect_multiscale_cv2(cc, frames[n % len(frames)], sf, 2, ms)
        n += 1
        if time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    return n / dt, n, dt, cv2.getNumThreads()


def run_reference(args, rank, world):
    """--impl reference: OpenCV CPU path on the host cores, rank 0 only."""
    if rank != 0:
        return
    from oracle import cv2_ref
    ge.load_package()                                       # only for the synthetic-frame generator; the .so is not loaded
    synth = __import__("importlib").import_module(ge.PKG_NAME + ".synth")
    w, h = map(int, args.size.split("x"))
    per_step = 8                                            # bounded sample of the workload per step
    sf, ms = args.scale_factor, (args.min_size, args.min_size)
    frames = make_frames(synth, w, h, per_step, seed0=0)
    import cv2
    cv2.setNumThreads(os.cpu_count() or 1)
    cc = cv2_ref.load(cv2_ref.cascade_path(args.cascade))
    for _ in range(args.warmup):
        cv2_ref.detect_multiscale_cv2(cc, frames[0], sf, 2, ms)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        for f in frames:
            cv2_ref.detect_multiscale_cv2(cc, f, sf, 2, ms)
    dt = time.perf_counter() - t0
    fps = args.steps * per_step / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8/int32/f32", "data": "synthetic",
        "config": {"workload": f"{os.path.basename(cv2_ref.cascade_path(args.cascade))} on synthetic {w}x{h} gray frames, "
                               f"scale {sf}, minNeighbors 2, minSize {ms[0]}x{ms[1]}", "frames_per_step": per_step},
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cv2.getNumThreads(), "kind": "reference",
                         "sample": f"{per_step} frames/step x {args.steps} steps, cv2 {cv2.__version__} "
                                   f"detectMultiScale, {os.cpu_count()} host cpus"},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


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
    synth = __import__("importlib").import_module(ge.PKG_NAME + ".synth")
    from oracle import cv2_ref

    w, h = map(int, args.size.split("x"))
    nF = args.frames
    casc = ug.Cascade(cv2_ref.cascade_path(args.cascade))
    GROUP = min(args.group, nF)
    sf, ms = args.scale_factor, (args.min_size, args.min_size)
    prm = ug.default_params(dense_stages=args.dense_stages, scale_factor=sf, min_size=ms)
    scan = tuple(map(int, args.scan_size.split("x"))) if args.scan_size else None
    det = ug.Detector(casc, prm, device=local, max_size=scan or (w, h), max_batch=GROUP)
    stream = torch.cuda.current_stream()
    det.set_stream(stream.cuda_stream)

    # rank r takes its own contiguous frame range (weak scaling: nF frames per rank per step)
    if scan:
        base = np.stack([synth.bgr_frame(w, h, synth.default_nfaces(w, h), 1000 * rank + i) for i in range(min(DISTINCT, nF))])
        frames_np = np.ascontiguousarray(np.concatenate([base] * ((nF + len(base) - 1) // len(base)))[:nF])
    else:
        frames_np = make_frames(synth, w, h, nF, seed0=1000 * rank)
    host = torch.from_numpy(frames_np).pin_memory()
    dev = host.cuda(non_blocking=False)
    host_np = host.numpy()
    sharding = __import__("importlib").import_module(ge.PKG_NAME + ".sharding")

    def gather(rects):
        """per-frame detection lists -> rank 0 (NCCL over NVLink), as fixed slots {n; rect[CAP]}"""
        if world == 1:
            return
        t = torch.from_numpy(sharding.pack_slots(rects, CAP)).cuda(non_blocking=True)
        sharding.gather_to_rank0(t, world, rank, nF)

    def step_device():
        if scan:
            r = det.detect_batch_scaled_ptr(dev.data_ptr(), True, nF, w, h, scan, cap=CAP)
        else:
            r = det.detect_batch_device(dev.data_ptr(), False, nF, w, h, cap=CAP)
        gather(r)
        return r

    def step_host():
        r = det.detect_batch_scaled(host_np, scan, cap=CAP) if scan else det.detect_batch(host_np, cap=CAP)
        gather(r)
        return r

    def timed(fn, steps, warmup, sample_clocks=False):
        sampler = ClockSampler(local) if sample_clocks and rank == 0 else None   # nvidia-smi needs ~1 s to start
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if sampler:
            sampler.t0 = time.time()
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        torch.cuda.synchronize()
        if sampler:
            sampler.t1 = time.time()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        clocks = sampler.stop() if sampler else None
        if world > 1:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, clocks

    ms_dev, clocks = timed(step_device, args.steps, max(args.warmup, 3), sample_clocks=True)
    ms_host, _ = timed(step_host, args.steps, 1)
    ctr = det.counters()                                   # counters of the last step (nF frames)

    # per-kernel device time (CUDA events on the launch stream, inside the library) -> roofline
    det.set_profiling(True)
    ktimes = {k: [0.0, 0] for k in ug.K_NAMES}
    for _ in range(max(2, min(args.steps, 3))):
        step_device()
        for k, (t, l) in det.kernel_times().items():
            ktimes[k][0] += t; ktimes[k][1] += l
    det.set_profiling(False)

    if rank == 0:
        hbm_peak, sm_max, peak_src = peaks()
        total_frames = nF * world * args.steps
        fps = total_frames / (ms_dev * 1e-3)
        fps_e2e = total_frames / (ms_host * 1e-3)
        plan_levels = ctr["levels"]; level_px = ctr["level_pixels"]
        pw, ph = scan or (w, h)                                 # size the pyramid is built from
        src_px = pw * ph
        sc, lw, lh, stp = det.plan_info(pw, ph)
        int_entries = int(sum((int(a) + 1) * (int(b) + 1) for a, b in zip(lw, lh)))
        weak_per_frame = ctr["weak_evals"] / max(ctr["frames"], 1)
        win_per_frame = ctr["windows_evaluated"] / max(ctr["frames"], 1)
        # algorithmic bytes per frame (DESIGN.md "Kernels and rooflines")
        alg = {
            "k_resize": src_px + level_px,                         # 1 B/source px read once + 1 B/level px written
            "k_colsum": level_px,                                  # 1 B/level px read (carry rows are noise)
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
            per_launch_ms = t / l
            kern[k] = {"ms_per_launch": per_launch_ms, "launches": l}
        total_k = sum(v["ms_per_launch"] for v in kern.values())
        for k, v in kern.items():
            v["share"] = v["ms_per_launch"] / total_k if total_k else 0.0
            if alg.get(k):
                gbs = alg[k] * GROUP / (v["ms_per_launch"] * 1e-3) / 1e9
                v.update({"bound": "hbm", "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak})
        casc_ms = kern["k_scan"]["ms_per_launch"] if "k_scan" in kern else 0.0   # both scan launches of a group
        l1_peak = 148 * 128 * sm_max * 1e6 / 1e9               # GB/s: 148 SMs x 128 B/clk x max SM clock
        casc_gbs = gather_bytes * GROUP / (casc_ms * 1e-3) / 1e9 if casc_ms else 0.0
        dominant = max(kern, key=lambda k: kern[k]["ms_per_launch"]) if kern else None
        if dominant == "k_scan":
            roof = {"kernel": "k_scan (window scan + all cascade stages)", "bound": "l1", "achieved": casc_gbs, "peak": l1_peak,
                    "unit": "GB/s", "frac": casc_gbs / l1_peak,
                    # ncu --set full, r01 (profiles/r01_ncu_k_scan_v7_summary.txt, the two launches of one 64-frame 720p
                    # launch group): dram read 348.1 + 615.5 MB, write 4.1 + 4.5 MB; the algorithmic HBM bytes are the
                    # S and SQ planes read once
                    "traffic": 972.2e6 if (w, h, GROUP) == (1280, 720, 64) else None,
                    "hbm": {"algorithmic_bytes": 8 * int_entries * GROUP,
                            "achieved": 8 * int_entries * GROUP / (casc_ms * 1e-3) / 1e9 if casc_ms else None,
                            "peak": hbm_peak, "unit": "GB/s",
                            "frac": 8 * int_entries * GROUP / (casc_ms * 1e-3) / 1e9 / hbm_peak if casc_ms else None},
                    "note": "gather/threshold work: bounded by L1/shared-memory throughput (148 SM x 128 B/clk x "
                            f"{sm_max:.0f} MHz), not HBM; algorithmic bytes = 4 B x (8 norm corners per window + "
                            "7.5 corners per weak classifier); the `hbm` sub-object gives the same launch against the "
                            "HBM roofline (planes read once) to show how far from HBM-bound it is"}
        else:
            v = kern.get(dominant, {})
            roof = {"kernel": dominant, "bound": v.get("bound", "hbm"), "achieved": v.get("achieved"), "peak": hbm_peak,
                    "unit": "GB/s", "frac": v.get("frac"), "traffic": None}
        hi = kern.get("k_integral", {})
        roof_hbm = {"kernel": "k_integral_bands (main pass of the integral)", "bound": "hbm", "achieved": hi.get("achieved"), "peak": hbm_peak,
                    "unit": "GB/s", "frac": hi.get("frac"),
                    # ncu r01 (profiles/r01_ncu_resize_integral_bands_summary.txt): dram read 191.1 MB + write 1030.5 MB per launch
                    "traffic": 1221.6e6 if (w, h, GROUP) == (1280, 720, 64) else None, "peak_source": peak_src}

        cpu = None
        if not args.no_cpu_baseline:
            if scan:                                            # reference pipeline: INTER_AREA resize + detectFaces (skin mask path)
                import cv2
                cv2.setNumThreads(os.cpu_count() or 1)
                cc = cv2_ref.load(cv2_ref.cascade_path(args.cascade))
                n_, t0_ = 0, time.perf_counter()
                while n_ < 128 and time.perf_counter() - t0_ < 12.0:
                    small = cv2.resize(host_np[n_ % DISTINCT], scan, interpolation=cv2.INTER_AREA)
                    cv2_ref.detect_faces_cv2(cc, small, sf)
                    n_ += 1
                cdt = time.perf_counter() - t0_
                cfps, cn, cthreads = n_ / cdt, n_, cv2.getNumThreads()
            else:
                cfps, cn, cdt, cthreads = cpu_reference_run(cv2_ref, host_np[:DISTINCT], args.cascade, 12.0, 128, sf, ms)
            cpu = {"value": cfps, "unit": "frames/s", "cores": cthreads, "kind": "reference",
                   "sample": f"{cn} of the step's {nF} frames in {cdt:.1f} s, cv2 detectMultiScale, "
                             f"{os.cpu_count()} host cpus"}
        groups = (nF + GROUP - 1) // GROUP
        launches_per_step = groups * (9 + (3 if scan else 0))   # resize, colsum, carry, integral, scan x2 (grid step 2 / step 1 levels), group init/link/finish (+ area, preprocess, equalize on the BGR path)
        line = {
            "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8/int32/f32", "data": "synthetic",
            "config": {"workload": (f"tracker front end: synthetic {w}x{h} BGR frames -> INTER_AREA to {scan[0]}x{scan[1]} -> detectFaces "
                                    f"(skin mask), {os.path.basename(cv2_ref.cascade_path(args.cascade))}, " if scan else
                                    f"{os.path.basename(cv2_ref.cascade_path(args.cascade))} on synthetic {w}x{h} gray frames, ")
                                   + f"scale {sf}, minNeighbors 2, minSize {ms[0]}x{ms[1]}"
                                   + (" (BASELINE configs[0]/[1] frame+params)" if (w, h, sf, ms[0]) == (1280, 720, 1.1, 30) and not scan else ""),
                       "frames_per_step_per_gpu": nF, "launch_group": GROUP, "levels": plan_levels,
                       "l2": f"inputs larger than L2: {nF * w * h / 1e6:.0f} MB of frames per step, "
                             f"{GROUP * (level_px + 8 * int_entries) / 1e6:.0f} MB of integrals per launch group",
                       "parallelism": f"frames sharded, {world} rank(s), lists gathered to rank 0"},
            "windows_evaluated_per_s": win_per_frame * fps,
            "weak_classifiers_per_s": weak_per_frame * fps,
            "e2e": {"value": fps_e2e, "unit": "frames/s", "h2d_bytes_per_step": nF * w * h * (3 if scan else 1),
                    "d2h_bytes_per_step": nF * (CAP * 16 + 4 + 64), "ms_per_step": ms_host / args.steps},
            "gpu_launches": launches_per_step * args.steps,
            "roofline": roof, "roofline_hbm": roof_hbm, "kernels": kern,
            "cpu_baseline": cpu, "clocks": clocks,
            "counters_last_step": ctr,
        }
        print(json.dumps(line), flush=True)
    det.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
