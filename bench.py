#!/usr/bin/env python
"""bench.py — ORB extraction throughput of the B200-native front end (frames/s), per the round contract.

A step = one pass of the hot path (pyramid -> FAST -> quadtree -> orientation/descriptors) over one batch of synthetic
frames.  The headline workload, at every N, is BASELINE.json configs[2]: 1920x1080 frames, nFeatures 2000, sharded by
frame over the ranks with no data-path collective (weak scaling).  `value` is measured with the batch resident in HBM,
`e2e` through the C-ABI with pinned HOST buffers (H2D of the frames and D2H of keypoints + descriptors inside the timed
region).  The line also carries, at N = 1, a `configs` array with the same figures for configs[0] (640x480 / 1000, plus
the single-frame host-to-host latency) and configs[1] (1280x720 / 1500); `extras` holds config 4 (2000 x 2000 Hamming)
with its CPU leg and issue-rate ceiling; at N > 1 `sweep` runs config 5 (125 frames per rank, NCCL all-gather of the
descriptor slabs, every cross-frame pair matched) and checks a sample of the pairs against a single-rank re-run.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl ours|reference]
"""
import argparse
import csv
import glob
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    "vga": dict(rows=480, cols=640, nfeat=1000, batch=512, config=0,
                label="640x480 synthetic frames, nFeatures 1000, scale 1.2, 8 levels, FAST 20/7 (BASELINE.json configs[0])"),
    "720p": dict(rows=720, cols=1280, nfeat=1500, batch=256, config=1,
                 label="1280x720 synthetic stream, nFeatures 1500, scale 1.2, 8 levels, FAST 20/7 (BASELINE.json configs[1])"),
    "1080p": dict(rows=1080, cols=1920, nfeat=2000, batch=256, config=2,
                  label="1920x1080 synthetic stream, nFeatures 2000, scale 1.2, 8 levels, FAST 20/7 (BASELINE.json configs[2])"),
}
HEADLINE = "1080p"
SWEEP_FRAMES_PER_RANK = 125          # config 5: 1000 frames on 8 GPUs
KERNELS = ["pyramid", "fast", "quadtree", "describe"]


def params(w):
    return (w["nfeat"], 1.2, 8, 20, 7)


def level_sizes(rows, cols, nlevels=8, sf=1.2):
    s = np.float32(1.0); out = []
    for l in range(nlevels):
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float32(cols) * inv)), int(np.rint(np.float32(rows) * inv))))
        s = np.float32(float(s) * float(np.float32(sf)))
    return out


def algorithmic_bytes(rows, cols, nkp):
    """bytes every stage must move when it reads its input once and writes its output once."""
    px = [w * h for w, h in level_sizes(rows, cols)]
    P = sum(px)
    return {"pyramid": (P - px[-1]) + (P - px[0]), "fast": P, "quadtree": 0, "describe": (961 + 1369 + 32 + 60) * nkp,
            "blur": 2 * P, "total": (P - px[-1]) + (P - px[0]) + P + 2 * P + (961 + 1369 + 32 + 60) * nkp}


class ClockSampler(threading.Thread):
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index; self.rows = []; self.stop = False; self.proc = None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits",
                                          "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")])
                if self.stop:
                    break
        except Exception:
            pass

    def finish(self):
        self.stop = True
        if self.proc:
            self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            for i, nme in enumerate(names):
                if len(r) > 3 + i and r[3 + i].lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback 6650 GB/s (B200_PROFILING.md)"


def ncu_traffic(workload, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per FRAME of `kernel` from the newest `ncu --set full` capture of this
    workload under profiles/ (profiles/index.json names the captures and the frames per launch of each)."""
    try:
        idx = json.load(open(os.path.join(ROOT, "profiles", "index.json")))
        caps = [c for c in idx["captures"] if c["workload"] == workload]
        if not caps:
            return None, None
        cap = caps[-1]
        rows = list(csv.reader(open(os.path.join(ROOT, "profiles", cap["raw_csv"]))))
        head, units = rows[0], rows[1]
        ik = head.index("Kernel Name"); ir = head.index("dram__bytes_read.sum"); iw = head.index("dram__bytes_write.sum")
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        for r in rows[2:]:
            if kernel + "_kernel" in r[ik]:
                tot = float(r[ir]) * scale[units[ir]] + float(r[iw]) * scale[units[iw]]
                return tot / cap["frames_per_launch"], "profiles/%s (ncu --set full, %d frames per launch)" % (cap["raw_csv"], cap["frames_per_launch"])
    except Exception:
        pass
    return None, None


def ncu_pipes(workload, kernel):
    """What the dominant kernel is really held by: busy fractions of the issue slots and of the integer pipes from the same
    committed `ncu --set full` capture the traffic figure comes from (counters, not a live measurement)."""
    try:
        idx = json.load(open(os.path.join(ROOT, "profiles", "index.json")))
        cap = [c for c in idx["captures"] if c["workload"] == workload][-1]
        rows = list(csv.reader(open(os.path.join(ROOT, "profiles", cap["raw_csv"]))))
        head = rows[0]
        ik = head.index("Kernel Name")
        names = {"issue_slots_busy_pct": "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
                 "alu_pipe_busy_pct": "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed",
                 "fma_heavy_pipe_busy_pct": "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
                 "dram_busy_pct": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"}
        for r in rows[2:]:
            if kernel + "_kernel" in r[ik]:
                out = {k: float(r[head.index(v)]) for k, v in names.items() if v in head}
                out["source"] = "profiles/%s" % cap["raw_csv"]
                return out
    except Exception:
        pass
    return None


# ------------------------------------------------------------------------------------------------------------------
# CPU legs (the only code in this file that touches oracle/)
def _reference_runner(w, threads):
    """Frame-parallel run of the reference's own ORBextractor.cc (oracle/_ref) — or, where that was not built, of the
    oracle port — over a bounded sample of the workload; returns (kind, nframes, run())."""
    from oracle import pyoracle
    from os1_b200 import synth
    nframes = max(threads, 8)
    frames = synth.video(nframes, w["rows"], w["cols"], seed=0)
    if pyoracle.Reference.available(canon=False):
        exts = [pyoracle.Reference(*params(w), canon=False) for _ in range(threads)]

        def run():
            nxt = [0]; lock = threading.Lock()

            def work(e):
                while True:
                    with lock:
                        f = nxt[0]; nxt[0] += 1
                    if f >= nframes:
                        return
                    e.extract(frames[f], cap=4096)
            ts = [threading.Thread(target=work, args=(e,)) for e in exts]
            [t.start() for t in ts]; [t.join() for t in ts]
        return "reference", nframes, run, frames

    def run():
        pyoracle.Oracle.extract_batch(frames, params(w), threads)
    return "port", nframes, run, frames


def opencv_check(w, frames):
    """how much slower the shim-backed reference's primitives are than real OpenCV.  One host thread.
    The reference's cv::resize / cv::GaussianBlur / cv::FAST calls (ORBextractor.cc:984,:950,:848,:854) go through the
    scalar restatements of oracle/cv_exact.h in oracle/_ref; here the same calls are timed on the same pyramid levels
    through those restatements and through cv2 (SIMD), and the reference's per-frame time is re-priced with cv2's."""
    try:
        import cv2
    except Exception as ex:
        return {"unavailable": "cv2 not importable here: %s" % ex}
    from oracle import pyoracle
    cv2.setNumThreads(1)
    img = frames[0]
    sizes = level_sizes(w["rows"], w["cols"])
    lv = [img]
    for (ww, hh) in sizes[1:]:
        lv.append(cv2.resize(lv[-1], (ww, hh), interpolation=cv2.INTER_LINEAR))

    def tmin(fn, reps=3):
        best = 1e9
        for _ in range(reps):
            t = time.perf_counter(); fn(); best = min(best, time.perf_counter() - t)
        return best * 1e3
    # the shim's primitives through the checker's C entry points with buffers allocated once (so that only the arithmetic is timed)
    L = pyoracle.Oracle.lib()
    P_ = pyoracle._p
    outs = [np.zeros((hh, ww), np.uint8) for (ww, hh) in sizes]
    fbuf = np.zeros((lv[0].size // 2 + 16, 3), np.int32)

    def shim_resize():
        for l in range(1, 8):
            L.orc_resize(P_(lv[l - 1]), lv[l - 1].shape[1], lv[l - 1].shape[0], lv[l - 1].strides[0], P_(outs[l]), sizes[l][0], sizes[l][1], sizes[l][0])

    def shim_blur():
        for l in range(8):
            L.orc_blur(P_(lv[l]), lv[l].shape[1], lv[l].shape[0], lv[l].strides[0], P_(outs[l]), outs[l].strides[0])

    def shim_fast(t):
        for a in lv:
            L.orc_fast(P_(a), a.shape[1], a.shape[0], a.strides[0], t, 1, P_(fbuf), len(fbuf))
    f20 = cv2.FastFeatureDetector_create(20, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    f7 = cv2.FastFeatureDetector_create(7, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    shim = {"pyramid": tmin(shim_resize), "blur": tmin(shim_blur), "fast20": tmin(lambda: shim_fast(20), 2), "fast7": tmin(lambda: shim_fast(7), 2)}
    real = {"pyramid": tmin(lambda: [cv2.resize(lv[l - 1], sizes[l], interpolation=cv2.INTER_LINEAR) for l in range(1, 8)]),
            "blur": tmin(lambda: [cv2.GaussianBlur(a, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101) for a in lv]),
            "fast20": tmin(lambda: [f20.detect(a) for a in lv], 2), "fast7": tmin(lambda: [f7.detect(a) for a in lv], 2)}
    ext = pyoracle.Reference(*params(w), canon=False) if pyoracle.Reference.available(canon=False) else pyoracle.Oracle(*params(w))
    whole = tmin(lambda: ext.extract(img, cap=4096), 3)
    # how many pixels the reference really hands to cv::FAST at each threshold (the per-cell windows overlap by 6 px; the retry at
    # minThFAST runs only in cells where iniThFAST found nothing): counted by the reference build's FAST hook
    P = float(sum(a.size for a in lv))
    w20, w7 = 1.0, 1.0 / 3.0
    if hasattr(ext, "fast_area"):
        ext.extract(img, cap=4096, record=True)
        a20, a7 = ext.fast_area()
        w20, w7 = a20 / P, a7 / P
    stage = lambda d: d["pyramid"] + d["blur"] + w20 * d["fast20"] + w7 * d["fast7"]
    # (the reference's FAST calls are per cell, a little less efficient than the whole-level calls timed here: the shim's share is
    # capped at the measured whole, so the re-priced time can never fall below what cv2's primitives alone take)
    rest = max(whole - stage(shim), 0.0)
    repriced = rest + stage(real)
    return {"threads": 1, "cv2": cv2.__version__, "reference_ms_per_frame": whole, "shim_stage_ms": shim, "cv2_stage_ms": real,
            "rest_ms_per_frame": rest,
            "repriced_ms_per_frame": repriced, "shim_slowdown_factor": whole / repriced,
            "fast_pixels_over_level_pixels": {"iniThFAST": w20, "minThFAST_retry": w7},
            "note": "whole-level calls of the same primitives on the same pyramid, FAST weighted by the pixels the reference really hands to "
                    "cv::FAST at each threshold; repriced = reference time with its three OpenCV primitives at cv2's speed"}


def cpu_baseline_sample(w, budget_s=10.0, with_opencv_check=True):
    threads = os.cpu_count() or 1
    kind, nframes, run, frames = _reference_runner(w, threads)
    run()
    reps = 0; t0 = time.perf_counter()
    while True:
        run(); reps += 1
        if time.perf_counter() - t0 > budget_s or reps >= 400:
            break
    dt = time.perf_counter() - t0
    out = {"value": reps * nframes / dt, "unit": "frames/s", "cores": threads, "kind": kind,
           "sample": "%d x %d synthetic %dx%d frames, one extractor per thread, %d host threads, %.1f s (%s)" % (
               reps, nframes, w["cols"], w["rows"], threads, dt,
               "oracle/_ref: unmodified reference ORBextractor.cc" if kind == "reference" else "oracle port")}
    if with_opencv_check:
        try:
            out["opencv_check"] = opencv_check(w, frames)
            f = out["opencv_check"].get("shim_slowdown_factor")
            if f:
                out["value_repriced_with_cv2_primitives"] = out["value"] * f
        except Exception as ex:
            out["opencv_check"] = {"unavailable": str(ex)}
    return out


def run_reference(args, rank, world):
    """`--impl reference`: the reference's own CPU implementation of the path on the host cores, all host threads, on the
    headline workload; each step a bounded sample (one frame per thread)."""
    if rank != 0:
        return
    w = WORKLOADS[HEADLINE]
    threads = os.cpu_count() or 1
    kind, nframes, step, _ = _reference_runner(w, threads)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    fps = args.steps * nframes / dt
    sample = "%d frames per step on %d host threads, one extractor per thread (%s)" % (
        nframes, threads, "oracle/_ref: unmodified reference ORBextractor.cc" if kind == "reference" else "oracle port")
    line = {"impl": "reference", "metric": "ORB frames/sec", "value": fps, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": {"workload": w["label"]},
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------------------
class Bench:
    """One workload on this rank's GPU: device-resident throughput, per-kernel split, end-to-end, H2D ceiling."""

    def __init__(self, torch, os1_b200, w, B, dev, rank, dist):
        from os1_b200 import synth
        self.torch, self.w, self.B, self.dev, self.dist = torch, w, B, dev, dist
        R, Cc = w["rows"], w["cols"]
        # every rank runs the SAME seeded stream: weak scaling means the same work per GPU, and FAST's cost follows the content
        # (with a stream per rank the step times of the ranks differ by up to 10 %; the step is the slowest rank's)
        self.frames_h = torch.from_numpy(synth.video(B, R, Cc, seed=1000)).pin_memory()
        self.frames_d = self.frames_h.to(dev, non_blocking=False)          # resident in HBM before any timed region
        self.ext = os1_b200.ORBextractor(*params(w), device=dev.index)
        self.cap = self.ext.max_keypoints()
        self.ext.reserve(R, Cc, B)
        self.xstream = torch.cuda.Stream(dev)                              # the extractor's stream: kernels and timing events share it
        self.ext.set_stream(self.xstream.cuda_stream)
        cap = self.cap
        self.kp_d = torch.empty((B, cap, 28), dtype=torch.uint8, device=dev)
        self.desc_d = torch.empty((B, cap, 32), dtype=torch.uint8, device=dev)
        self.cnt_d = torch.zeros(B, dtype=torch.int32, device=dev)
        self.kp_h = torch.empty((B, cap, 28), dtype=torch.uint8).pin_memory()
        self.desc_h = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
        self.cnt_h = torch.zeros(B, dtype=torch.int32).pin_memory()

    def step_device(self, n=None):
        w = self.w; n = n or self.B
        self.ext.extract_batch_raw(self.frames_d.data_ptr(), n, w["rows"], w["cols"], w["cols"], w["rows"] * w["cols"], True,
                                   self.kp_d.data_ptr(), self.desc_d.data_ptr(), self.cap, self.cnt_d.data_ptr(), True)

    def step_e2e(self):
        w = self.w
        self.ext.extract_batch_raw(self.frames_h.data_ptr(), self.B, w["rows"], w["cols"], w["cols"], w["rows"] * w["cols"], False,
                                   self.kp_h.data_ptr(), self.desc_h.data_ptr(), self.cap, self.cnt_h.data_ptr(), False)

    def barrier(self):
        self.torch.cuda.synchronize(self.dev)
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize(self.dev)

    def reduce_max(self, v):
        if self.dist is None:
            return v
        t = self.torch.tensor([v], device=self.dev, dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def measure_device(self, steps, warmup, sampler=None):
        """K steps enqueued back to back between two events on the extractor's stream — no host synchronisation inside."""
        torch = self.torch
        for _ in range(warmup):
            self.step_device()
        self.ext.synchronize()
        pre_rows = None
        if sampler is not None:
            t_wait = time.time()
            while len(sampler.rows) < 2 and time.time() - t_wait < 3.0:    # keep the GPU busy until the sampler is reporting
                self.step_device(); self.ext.synchronize()
            pre_rows = list(sampler.rows[-2:])
            sampler.rows = []
        self.barrier()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(self.xstream)
        for _ in range(steps):
            self.step_device()
        e1.record(self.xstream)
        self.barrier()
        ms_local = e0.elapsed_time(e1)
        ms = self.reduce_max(ms_local)
        self.ext.synchronize()                                              # (reads the device-side status word: capacity errors surface here)
        clocks = None
        if sampler is not None:
            if not sampler.rows:
                sampler.rows = pre_rows or []
                src = "under load immediately before the timed region (it was shorter than the 50 ms sampling period)"
            else:
                src = "during the timed region"
            pw = [float(r[2]) for r in sampler.rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
            clocks = sampler.finish(); clocks["sampled"] = src
            if self.dist is not None:
                # the step is the slowest rank's: every rank's own time, clock and power, so that a slow GPU shows as such
                mine = torch.tensor([ms_local / steps, clocks["sm_mhz"] or 0.0, float(np.median(pw)) if pw else 0.0], device=self.dev, dtype=torch.float64)
                allr = torch.empty((self.dist.get_world_size(), 3), device=self.dev, dtype=torch.float64)
                self.dist.all_gather_into_tensor(allr.view(-1), mine)
                a = allr.cpu().numpy()
                clocks["per_rank"] = {"ms_per_step": [round(float(v), 4) for v in a[:, 0]], "sm_mhz": [float(v) for v in a[:, 1]],
                                      "power_w": [round(float(v), 1) for v in a[:, 2]]}
        return ms / steps, clocks

    def measure_stages(self, reps=5):
        """Per-kernel durations (CUDA events around each launch), in a separate untimed pass."""
        self.ext.set_timing(True)
        acc = np.zeros(4)
        for _ in range(reps):
            self.step_device()
            acc += self.ext.stage_times_ms()
        self.ext.set_timing(False)
        return acc / reps

    def measure_e2e(self, steps, warmup):
        for _ in range(max(1, warmup // 2)):
            self.step_e2e()
        self.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            self.step_e2e()
        self.torch.cuda.synchronize(self.dev)
        s = self.reduce_max(time.perf_counter() - t0)
        return s / steps * 1e3                 # ms per step

    def measure_e2e_two_handles(self, steps):
        """The same host-to-host call from two host threads with a handle each (ctypes drops the GIL inside the call): while one
        handle drains the last chunks of its batch the other is already uploading, so the link does not idle between batches.
        Reported beside the single-handle figure, which stays the headline."""
        import threading
        import os1_b200
        torch, w, B, cap = self.torch, self.w, self.B, self.cap
        ext2 = os1_b200.ORBextractor(*params(w), device=self.dev.index)
        ext2.reserve(w["rows"], w["cols"], B)
        kp2 = torch.empty((B, cap, 28), dtype=torch.uint8).pin_memory(); d2 = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
        c2 = torch.zeros(B, dtype=torch.int32).pin_memory()

        def step2():
            ext2.extract_batch_raw(self.frames_h.data_ptr(), B, w["rows"], w["cols"], w["cols"], w["rows"] * w["cols"], False,
                                   kp2.data_ptr(), d2.data_ptr(), cap, c2.data_ptr(), False)
        for _ in range(2):
            step2()
        n = max(1, steps // 2)
        th = [threading.Thread(target=lambda f=f: [f() for _ in range(n)]) for f in (self.step_e2e, step2)]
        torch.cuda.synchronize(self.dev)
        t0 = time.perf_counter()
        for t in th:
            t.start()
        for t in th:
            t.join()
        torch.cuda.synchronize(self.dev)
        ms = (time.perf_counter() - t0) / (2 * n) * 1e3
        assert np.array_equal(c2.numpy(), self.cnt_h.numpy()), "the two handles disagree"
        for f in (0, B - 1):
            assert np.array_equal(d2[f, :int(c2[f])].numpy(), self.desc_h[f, :int(c2[f])].numpy()), "the two handles disagree"
        ext2.close()
        return ms

    def measure_h2d_ceiling(self, steps):
        """The same frame bytes from the same pinned buffer into the same device buffer with plain asynchronous copies and
        no kernels, all ranks at once: what the host->device feed allows."""
        torch = self.torch
        scratch = torch.empty_like(self.frames_d)
        st = torch.cuda.Stream(self.dev)
        with torch.cuda.stream(st):
            scratch.copy_(self.frames_h, non_blocking=True)
        self.barrier()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(st):
            e0.record(st)
            for _ in range(steps):
                scratch.copy_(self.frames_h, non_blocking=True)
            e1.record(st)
        self.barrier()
        ms = self.reduce_max(e0.elapsed_time(e1)) / steps
        del scratch
        return ms

    def record(self, steps, warmup, world, sampler=None, with_ceiling=True):
        w, B = self.w, self.B
        ms, clocks = self.measure_device(steps, warmup, sampler)
        launches_per_step = int(self.ext.launches_per_call())            # kernels the last device-resident step launched (counted by the library)
        stage_ms = self.measure_stages()
        counts = self.cnt_d.cpu().numpy()
        e2e_ms = self.measure_e2e(steps, warmup)
        assert np.array_equal(self.cnt_h.numpy(), counts), "host and device paths disagree on keypoint counts"
        h2d_ms = self.measure_h2d_ceiling(max(3, min(steps, 20))) if with_ceiling else None
        nkp = float(counts.sum()) / B
        alg = algorithmic_bytes(w["rows"], w["cols"], nkp)
        dom = int(np.argmax(stage_ms))
        peak, which = peaks()
        dom_bytes = alg[KERNELS[dom]] * B
        achieved = dom_bytes / (stage_ms[dom] * 1e-3) / 1e9
        whole = alg["total"] * B / (ms * 1e-3) / 1e9
        traffic_pf, traffic_src = ncu_traffic([k for k, v in WORKLOADS.items() if v is w][0], KERNELS[dom])
        per_kernel = {}
        for i, k in enumerate(KERNELS):
            if alg[k]:
                g = alg[k] * B / (stage_ms[i] * 1e-3) / 1e9
                per_kernel[k] = {"ms_per_launch": float(stage_ms[i]), "achieved_gbs": g, "frac": g / peak}
            else:
                per_kernel[k] = {"ms_per_launch": float(stage_ms[i])}
        roof = {"bound": "hbm", "kernel": KERNELS[dom] + "_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic_pf * B if traffic_pf else None, "traffic_source": traffic_src, "peak_source": which,
                "algorithmic_bytes_per_launch": dom_bytes, "kernels": per_kernel,
                "kernel_sum_ms": float(stage_ms.sum()), "kernel_sum_over_step": float(stage_ms.sum() / ms),
                "whole_step_achieved_gbs": whole, "whole_step_frac": whole / peak,
                "pipes": ncu_pipes([k for k, v in WORKLOADS.items() if v is w][0], KERNELS[dom]),
                "note": "FAST is bound by the integer ALU pipe and instruction issue, not by HBM: see DESIGN.md section 4 "
                        "(`pipes`: busy fractions from the committed ncu capture)"}
        h2d_bytes = int(B * w["rows"] * w["cols"])
        rec = {"value": world * B / (ms * 1e-3), "unit": "frames/s", "ms_per_step": ms,
               "config": {"workload": w["label"], "frames_per_step_per_gpu": B, "keypoints_per_frame": nkp,
                          "l2": "inputs larger than L2 (%d MB of frames per step per GPU, plus the pyramid)" % (h2d_bytes >> 20),
                          "parallelism": "frame-sharded, %d rank(s), no data-path collective; every rank runs the same seeded stream" % world},
               "e2e": {"value": world * B / (e2e_ms * 1e-3), "unit": "frames/s", "h2d_bytes_per_step": h2d_bytes,
                       "d2h_bytes_per_step": int(B * self.cap * 60 + B * 4 + 4), "ms_per_step": e2e_ms},
               "roofline": roof, "launches_per_step": launches_per_step}
        if with_ceiling and world == 1:
            ms2 = self.measure_e2e_two_handles(steps)
            rec["e2e"]["two_handles"] = {"value": B / (ms2 * 1e-3), "unit": "frames/s", "ms_per_step": ms2,
                                         "how": "the same call from two host threads with a handle each, batches alternating"}
        if h2d_ms:
            ceil_fps = world * B / (h2d_ms * 1e-3)
            rec["h2d_ceiling"] = {"frames_per_s": ceil_fps, "aggregate_gbs": world * h2d_bytes / (h2d_ms * 1e-3) / 1e9, "ms_per_step": h2d_ms,
                                  "e2e_frac_of_ceiling": rec["e2e"]["value"] / ceil_fps,
                                  "how": "plain cudaMemcpyAsync of the step's frames from the same pinned buffer, all ranks at once, no kernels"}
        if clocks is not None:
            rec["clocks"] = clocks
        return rec

    def close(self):
        self.ext.close()


def latency_record(os1_b200, w, dev):
    """Single frame, host buffers in and out, through os1_orb_extract — what Frame.cc:133 sees (config 1's latency target: < 0.5 ms).
    The image is ordinary pageable memory, as a cv::Mat is; the output buffers are allocated once, as the C++ adapter does."""
    import ctypes as C
    from os1_b200 import synth, _lib
    from os1_b200.orb import KEYPOINT_DTYPE
    img = synth.frame(w["rows"], w["cols"], 0)
    e = os1_b200.ORBextractor(*params(w), device=dev.index)
    cap = e.max_keypoints()
    kp = np.zeros(cap, KEYPOINT_DTYPE); desc = np.zeros((cap, 32), np.uint8); n = C.c_int(0)
    L = _lib.load()
    call = lambda: L.os1_orb_extract(e._h, img.ctypes.data, img.shape[0], img.shape[1], img.strides[0], kp.ctypes.data, cap, desc.ctypes.data, C.byref(n))
    for _ in range(20):
        _lib.check(call(), e._h)
    ts = []
    for _ in range(300):
        t = time.perf_counter(); rc = call(); ts.append(time.perf_counter() - t)
        if rc:
            _lib.check(rc, e._h)
    e.close()
    return float(np.median(ts) * 1e3)


def issue_ceiling(dev_index):
    """POPC / LOP3 issue rates measured by the micro-kernels of os1_b200/csrc/ubench.cu (measure, do not assume)."""
    try:
        from os1_b200 import ubench
        return ubench.measure(dev_index)
    except Exception as ex:
        return {"unavailable": str(ex)}


def extras(torch, os1_b200, dev):
    """Config 4 (2000 x 2000 Hamming, with its CPU leg and issue ceiling), the one-GPU frame-pair sweep, and
    MapPoint::ComputeDistinctiveDescriptors."""
    from os1_b200 import synth, matcher
    out = {}
    ceil = issue_ceiling(dev.index)
    out["issue_rates"] = ceil
    peak_pairs = ceil.get("hamming_pairs_per_s_ceiling") if isinstance(ceil, dict) else None

    def hroof(gpairs):
        if not peak_pairs:
            return None
        return {"bound": "popc+lop3 issue", "peak": peak_pairs / 1e9, "achieved": gpairs, "unit": "G pairs/s", "frac": gpairs / (peak_pairs / 1e9)}
    # config 4: 2000 x 2000 descriptors of two frames extracted with the 2000-feature extractor, all on device
    fr = synth.video(2, 720, 1280, seed=9)
    e2 = os1_b200.ORBextractor(2000, 1.2, 8, 20, 7, device=dev.index)
    d0 = e2(fr[0])[1][:2000]; d1 = e2(fr[1])[1][:2000]
    e2.close()
    a = torch.from_numpy(np.ascontiguousarray(d0)).to(dev); b = torch.from_numpy(np.ascontiguousarray(d1)).to(dev)
    o = [torch.zeros(len(d0), dtype=torch.int32, device=dev) for _ in range(3)]
    s = torch.cuda.current_stream(dev)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    pairs = len(d0) * len(d1)
    reps = 200
    for name, fn in (("hamming_knn2", matcher.hamming_knn2_raw), ("hamming_knn2_noalloc", matcher.hamming_knn2_noalloc_raw)):
        call = lambda: fn(a.data_ptr(), len(d0), b.data_ptr(), len(d1), o[0].data_ptr(), o[1].data_ptr(), o[2].data_ptr(), device=dev.index,
                          stream=s.cuda_stream)
        for _ in range(20):
            call()
        torch.cuda.synchronize(dev); e0.record(s)
        for _ in range(reps):
            call()
        e1.record(s); torch.cuda.synchronize(dev)
        us = e0.elapsed_time(e1) * 1e3 / reps
        rec = {"us_per_call": us, "gpairs_per_s": pairs / us / 1e3, "gops_per_s": 24 * pairs / us / 1e3,
               "note": "1 pair = 8 XOR + 8 POPC + 8 ADD on 32-bit words (SURVEY.md 8d); device pointers"}
        # the same call replayed from a CUDA graph (50 launches per graph): the GPU's own time per call, without the host's launch work
        try:
            gst = torch.cuda.Stream(dev)
            gcall = lambda: matcher.hamming_knn2_noalloc_raw(a.data_ptr(), len(d0), b.data_ptr(), len(d1), o[0].data_ptr(), o[1].data_ptr(), o[2].data_ptr(),
                                                             device=dev.index, stream=gst.cuda_stream)
            if name == "hamming_knn2_noalloc":
                gcall(); torch.cuda.synchronize(dev)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=gst):
                    for _ in range(50):
                        gcall()
                g.replay(); torch.cuda.synchronize(dev)
                e0.record(s)
                for _ in range(10):
                    g.replay()
                e1.record(s); torch.cuda.synchronize(dev)
                usg = e0.elapsed_time(e1) * 1e3 / 500
                rec["us_per_call_gpu_only"] = usg
                rec["gpairs_per_s_gpu_only"] = pairs / usg / 1e3
                rec["gpu_only_note"] = "50 calls captured in a CUDA graph and replayed: no host launch work between the kernels"
        except Exception as ex:
            rec["us_per_call_gpu_only"] = None; rec["gpu_only_note"] = "graph replay failed: %s" % ex
        r = hroof(rec.get("gpairs_per_s_gpu_only") or rec["gpairs_per_s"])
        if r:
            rec["roofline"] = r
        out["%s_%dx%d" % (name, len(d0), len(d1))] = rec
    # B4: the reference's SWAR loop (ORBmatcher.cc:1605-1621) as the same 2000 x 2000 best / second-best sweep on the host
    try:
        from oracle import pyoracle
        threads = os.cpu_count() or 1
        cpu = {}
        for nm, th in (("1_thread", 1), ("all_cores", threads)):
            pyoracle.Oracle.hamming_knn2(d0, d1, threads=th)
            n = 0; t0 = time.perf_counter()
            while time.perf_counter() - t0 < 1.5:
                ref = pyoracle.Oracle.hamming_knn2(d0, d1, threads=th); n += 1
            dt = (time.perf_counter() - t0) / n
            cpu[nm] = {"threads": th, "ms_per_call": dt * 1e3, "gpairs_per_s": pairs / dt / 1e9, "gops_per_s": 24 * pairs / dt / 1e9}
        cpu["agrees_with_gpu"] = bool(all(np.array_equal(x, y.cpu().numpy()) for x, y in zip(ref, o)))
        cpu["kind"] = "port (the reference's 8-word SWAR popcount, g++ -O3)"
        out["hamming_knn2_cpu_%dx%d" % (len(d0), len(d1))] = cpu
    except Exception as ex:
        out["hamming_knn2_cpu"] = {"unavailable": str(ex)}
    # cross-frame sweep on one GPU: all pairs of 48 frames of a video, one launch
    from os1_b200 import shard
    F = 48
    vid = synth.video(F, 720, 1280, seed=77)
    e3 = os1_b200.ORBextractor(1500, 1.2, 8, 20, 7, device=dev.index)
    kp_, desc_, cnt_ = e3.extract_batch(vid)
    e3.close()
    dsl = torch.from_numpy(desc_).to(dev); csl = torch.from_numpy(cnt_).to(dev)
    prs = shard.pair_partition(F, 0, 1)
    shard.sweep_pairs(dsl, csl, prs[:8], dev.index)
    torch.cuda.synchronize(dev); e0.record(s)
    acc = shard.sweep_pairs(dsl, csl, prs, dev.index)
    e1.record(s); torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1)
    dpairs = float((cnt_[prs[:, 0]].astype(np.float64) * cnt_[prs[:, 1]]).sum())
    rec = {"frame_pairs": int(len(prs)), "descriptor_pairs": dpairs, "ms": ms, "gpairs_per_s": dpairs / ms / 1e6,
           "gops_per_s": 24 * dpairs / ms / 1e6, "accepted_matches": int(acc.sum().item())}
    r = hroof(rec["gpairs_per_s"])
    if r:
        rec["roofline"] = r
    out["hamming_sweep_%d_frames" % F] = rec
    # MapPoint::ComputeDistinctiveDescriptors for a map's worth of points in one launch (device pointers), with the
    # reference's algorithm (oracle port, one host thread) on a bounded sample beside it
    from os1_b200 import _lib
    rng = np.random.default_rng(4)
    M = 50000
    cnt = rng.integers(2, 25, M); ptr = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int32)
    dd_h = rng.integers(0, 256, (int(ptr[-1]), 32), dtype=np.uint8)
    dd = torch.from_numpy(dd_h).to(dev); dp = torch.from_numpy(ptr).to(dev)
    bi = torch.empty(M, dtype=torch.int32, device=dev); bm = torch.empty_like(bi)
    calld = lambda: _lib.check(_lib.load().os1_distinctive_descriptors(dd.data_ptr(), dp.data_ptr(), M, bi.data_ptr(), bm.data_ptr(), 1, dev.index,
                                                                        s.cuda_stream))
    for _ in range(3):
        calld()
    torch.cuda.synchronize(dev); e0.record(s)
    for _ in range(10):
        calld()
    e1.record(s); torch.cuda.synchronize(dev)
    msd = e0.elapsed_time(e1) / 10
    rec = {"map_points": M, "observations": int(ptr[-1]), "ms_per_call": msd, "mpoints_per_s": M / msd / 1e3,
           "note": "2..24 observations per point; the call includes its status read-back"}
    try:
        from oracle import pyoracle              # cpu-baseline leg: the checker's port, timed beside the GPU figure
        Ms = 5000
        t = time.perf_counter(); obi, _ = pyoracle.Oracle.distinctive_descriptors(dd_h[:ptr[Ms]], ptr[:Ms + 1]); dt = time.perf_counter() - t
        rec["cpu_port_mpoints_per_s"] = Ms / dt / 1e6
        rec["cpu_port_sample"] = "%d points, one host thread, %.3f s" % (Ms, dt)
        rec["agrees_with_cpu_port"] = bool(np.array_equal(obi, bi[:Ms].cpu().numpy()))
    except Exception as ex:
        rec["cpu_port_sample"] = "unavailable: %s" % ex
    out["distinctive_descriptors"] = rec
    return out


def sweep_config5(torch, dist, hb, rank, world, local_rank):
    """BASELINE.json configs[4]: every rank extracts its 125 frames of the 1080p / 2000 stream, the descriptor slabs and counts are
    all-gathered over NCCL, the upper triangle of frame pairs is dealt block-cyclically to the ranks and matched in one
    batched launch per rank; a sample of the pairs is then re-run on rank 0 alone and must give identical counts."""
    from os1_b200 import shard
    dev = hb.dev
    Fl = min(SWEEP_FRAMES_PER_RANK, hb.B)
    # the sweep is over DISTINCT frames: every rank gets its own seeded stream here (the headline measurement is over)
    from os1_b200 import synth
    hb.frames_d[:Fl].copy_(torch.from_numpy(synth.video(Fl, hb.w["rows"], hb.w["cols"], seed=2000 + rank)))
    stream = torch.cuda.current_stream(dev)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    hb.step_device(Fl); hb.ext.synchronize()
    dsl = hb.desc_d[:Fl].contiguous(); csl = hb.cnt_d[:Fl].contiguous()
    shard.allgather_descriptors(dsl, csl)                         # warm-up (NCCL channel set-up)
    pairs = shard.pair_partition(Fl * world, rank, world)
    hb.barrier()
    # extraction of this rank's frames (on the extractor's stream), then the exchange and the match on the current stream
    x0 = torch.cuda.Event(enable_timing=True); x1 = torch.cuda.Event(enable_timing=True)
    x0.record(hb.xstream); hb.step_device(Fl); x1.record(hb.xstream)
    stream.wait_event(x1)
    ev[0].record(stream)
    dsl = hb.desc_d[:Fl].contiguous(); csl = hb.cnt_d[:Fl].contiguous()
    d_all, c_all = shard.allgather_descriptors(dsl, csl)
    ev[1].record(stream)
    acc = shard.sweep_pairs(d_all, c_all, pairs, local_rank)
    ev[2].record(stream)
    hb.barrier()
    counts = c_all.cpu().numpy().astype(np.float64)
    dpairs_local = float((counts[pairs[:, 0]] * counts[pairs[:, 1]]).sum()) if len(pairs) else 0.0
    tg = torch.tensor([x0.elapsed_time(x1), ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2])], device=dev, dtype=torch.float64)
    tmax = tg.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ts = torch.tensor([float(len(pairs)), dpairs_local, float(acc.sum().item())], device=dev, dtype=torch.float64)
    dist.all_reduce(ts, op=dist.ReduceOp.SUM)
    # correctness: gather every rank's per-pair accepted counts, re-run a sample of pairs on rank 0 alone
    npl = (Fl * world * (Fl * world - 1) // 2 + world - 1) // world
    pad = torch.full((npl,), -1, dtype=torch.int32, device=dev); pad[:len(pairs)] = acc
    allacc = torch.empty((world, npl), dtype=torch.int32, device=dev)
    dist.all_gather_into_tensor(allacc.view(-1), pad)
    check = None
    if rank == 0:
        gi, gj = np.triu_indices(Fl * world, k=1)
        rng = np.random.default_rng(5)
        nsample = min(len(gi), 2000)
        pick = np.sort(rng.choice(len(gi), nsample, replace=False))
        mine = shard.sweep_pairs(d_all, c_all, np.stack([gi[pick], gj[pick]], 1), local_rank).cpu().numpy()
        theirs = allacc.cpu().numpy()[pick % world, pick // world]
        check = {"pairs_rechecked_on_rank0": int(nsample), "identical": bool(np.array_equal(mine, theirs)),
                 "mismatching_pairs": int((mine != theirs).sum())}
        assert check["identical"], "sharded sweep disagrees with the single-rank re-run: %r" % check
    ext_ms, ag_ms, match_ms = (float(v) for v in tmax.tolist())
    npairs, dpairs, accepted = (float(v) for v in ts.tolist())
    return {"config": "BASELINE.json configs[4]: %d frames of 1920x1080 / 2000 features (%d per rank), all cross-frame pairs" % (Fl * world, Fl),
            "frames": Fl * world, "frame_pairs": int(npairs), "descriptor_pairs": dpairs, "extract_ms": ext_ms, "allgather_ms": ag_ms,
            "allgather_bytes_per_rank": int(dsl.numel() + csl.numel() * 4), "match_ms": match_ms, "total_ms": ext_ms + ag_ms + match_ms,
            "accepted_matches": int(accepted), "gpairs_per_s": dpairs / (match_ms * 1e-3) / 1e9, "gops_per_s": 24 * dpairs / (match_ms * 1e-3) / 1e9,
            "check": check,
            "note": "descriptor pairs/s over all ranks, one batched launch per rank incl. the acceptance test of ORBmatcher.cc:454-456; "
                    "NCCL all_gather_into_tensor; times are the max over ranks"}


# ------------------------------------------------------------------------------------------------------------------
def run_ours(args, rank, world, local_rank):
    import torch
    import os1_b200
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    w = WORKLOADS[HEADLINE]
    B = args.batch or w["batch"]
    hb = Bench(torch, os1_b200, w, B, dev, rank, dist)
    sampler = ClockSampler(local_rank); sampler.start()          # nvidia-smi needs ~0.3 s to come up: start before warm-up
    head = hb.record(args.steps, args.warmup, world, sampler)
    sweep = sweep_config5(torch, dist, hb, rank, world, local_rank) if dist is not None else None
    hb.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    line = {"metric": "ORB frames/sec", "value": head["value"], "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic", "config": head["config"], "e2e": head["e2e"], "gpu_launches": int(args.steps * head["launches_per_step"] * world),
            "roofline": head["roofline"], "clocks": head.get("clocks"), "h2d_ceiling": head.get("h2d_ceiling")}
    if sweep is not None:
        line["sweep"] = sweep
    if world == 1 and args.quick:
        line["note"] = "--quick: headline workload only (profiling runs); cpu_baseline, configs and extras skipped"
    if world == 1 and not args.quick:
        try:
            line["cpu_baseline"] = cpu_baseline_sample(w, 10.0)
        except Exception as ex:           # the checker libraries are optional on the box; say so rather than fail the bench
            line["cpu_baseline"] = {"value": None, "unit": "frames/s", "cores": 0, "kind": "port", "sample": "unavailable: %s" % ex}
        # the other two configurations of BASELINE.json's metric line, same figures
        subs = []
        for key in ("vga", "720p"):
            ww = WORKLOADS[key]
            try:
                sb = Bench(torch, os1_b200, ww, ww["batch"], dev, 0, None)
                rec = sb.record(max(5, min(args.steps, 30)), 3, 1, None)
                sb.close()
                rec["metric"] = "ORB frames/sec"; rec["n_gpus"] = 1; rec["baseline_config_index"] = ww["config"]
                if key == "vga":
                    rec["latency_ms_single_frame_host_to_host"] = latency_record(os1_b200, ww, dev)
                try:
                    rec["cpu_baseline"] = cpu_baseline_sample(ww, 6.0, with_opencv_check=False)
                except Exception as ex:
                    rec["cpu_baseline"] = {"value": None, "sample": "unavailable: %s" % ex}
                subs.append(rec)
            except Exception as ex:
                subs.append({"config": {"workload": ww["label"]}, "error": str(ex)})
        line["configs"] = subs
        try:
            line["latency_ms_1920x1080_2000feat_host_to_host"] = latency_record(os1_b200, w, dev)
        except Exception as ex:
            line["latency_ms_1920x1080_2000feat_host_to_host"] = None
        try:
            line["extras"] = extras(torch, os1_b200, dev)
        except Exception as ex:
            line["extras"] = {"error": str(ex)}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=0, help="frames per step per GPU (default: the workload's)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--quick", action="store_true", help="headline workload only: no CPU legs, sub-configs or extras (ncu captures)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
