#!/usr/bin/env python
"""bench.py — ORB extraction throughput of the B200-native front end (frames/s), per the round contract.

A step = one pass of the hot path (pyramid -> FAST -> quadtree -> orientation/descriptors) over one batch of
synthetic 1280x720 frames, 1500 features (BASELINE.json configs[1]); frames are sharded by rank with no
data-path collective (weak scaling).  `value` is measured with the batch resident in HBM, `e2e` through the
C-ABI with pinned HOST buffers (H2D of the frames and D2H of keypoints+descriptors inside the timed region).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl ours|reference]
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

ROWS, COLS, NFEAT = 720, 1280, 1500
PARAMS = (NFEAT, 1.2, 8, 20, 7)
WORKLOAD = "1280x720 synthetic stream, nFeatures 1500, scale 1.2, 8 levels, FAST 20/7 (BASELINE.json configs[1])"


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
        return json.load(open(p))["hbm_gbs"], "measured"
    return 6650.0, "fallback"


# ------------------------------------------------------------------------------------------------------------------
def run_reference(args, rank, world):
    """The reference's own CPU implementation of the path on the host cores: oracle/_ref (the unmodified
    ORBextractor.cc) when it was built, else the oracle port; one extractor per thread, all host threads."""
    if rank != 0:
        return
    from oracle import pyoracle
    from os1_b200 import synth
    threads = os.cpu_count() or 1
    nframes = max(threads, 8)                      # bounded sample per step
    frames = synth.video(nframes, ROWS, COLS, seed=0)
    use_ref = pyoracle.Reference.available(canon=False)
    if use_ref:
        exts = [pyoracle.Reference(*PARAMS, canon=False) for _ in range(threads)]
        kind = "reference"

        def step():
            nxt = [0]; lock = threading.Lock(); tot = [0]

            def work(e):
                while True:
                    with lock:
                        f = nxt[0]; nxt[0] += 1
                    if f >= nframes:
                        return
                    k, _ = e.extract(frames[f], cap=4096)
                    with lock:
                        tot[0] += len(k)
            ts = [threading.Thread(target=work, args=(e,)) for e in exts]
            [t.start() for t in ts]; [t.join() for t in ts]
            return tot[0]
    else:
        kind = "port"

        def step():
            tot, _ = pyoracle.Oracle.extract_batch(frames, PARAMS, threads)
            return tot
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    fps = args.steps * nframes / dt
    sample = "%d frames per step on %d host threads, one extractor per thread (%s)" % (
        nframes, threads, "oracle/_ref: unmodified reference ORBextractor.cc" if use_ref else "oracle port")
    line = {"impl": "reference", "metric": "ORB frames/sec", "value": fps, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": {"workload": WORKLOAD, "frames_per_step": nframes},
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def cpu_baseline_sample():
    """Bounded CPU sample for the `cpu_baseline` object of our own line (rank 0, N=1)."""
    from oracle import pyoracle
    from os1_b200 import synth
    threads = os.cpu_count() or 1
    nframes = max(threads, 8)
    frames = synth.video(nframes, ROWS, COLS, seed=0)
    if pyoracle.Reference.available(canon=False):
        kind = "reference"
        exts = [pyoracle.Reference(*PARAMS, canon=False) for _ in range(threads)]

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
    else:
        kind = "port"

        def run():
            pyoracle.Oracle.extract_batch(frames, PARAMS, threads)
    run()
    reps = 0; t0 = time.perf_counter()
    while True:
        run(); reps += 1
        if time.perf_counter() - t0 > 12.0 or reps >= 400:
            break
    dt = time.perf_counter() - t0
    return {"value": reps * nframes / dt, "unit": "frames/s", "cores": threads, "kind": kind,
            "sample": "%d x %d synthetic 720p frames, one extractor per thread, %d host threads, %.1f s" % (reps, nframes, threads, dt)}


def extras(torch, os1_b200, dev):
    """Secondary figures of BASELINE.json's metric line: single-frame latency (config 1) and Hamming throughput (config 4)."""
    from os1_b200 import synth, matcher
    out = {}
    # config 1: one 640x480 frame, 1000 features, host buffers in and out (what Frame.cc:133 sees)
    img = synth.frame(480, 640, 0)
    e = os1_b200.ORBextractor(1000, 1.2, 8, 20, 7, device=dev.index)
    for _ in range(20):
        e(img)
    ts = []
    for _ in range(200):
        t = time.perf_counter(); e(img); ts.append(time.perf_counter() - t)
    out["latency_ms_640x480_1000feat_host_to_host"] = float(np.median(ts) * 1e3)
    # the other two resolutions of BASELINE.json's metric line (configs[0] and configs[2]), frames resident in HBM
    for name, (hh, ww, nfeat, nb) in (("640x480_1000feat", (480, 640, 1000, 512)), ("1920x1080_2000feat", (1080, 1920, 2000, 128))):
        ex = os1_b200.ORBextractor(nfeat, 1.2, 8, 20, 7, device=dev.index)
        vid = torch.from_numpy(synth.video(nb, hh, ww, seed=31)).to(dev)
        capx = ex.max_keypoints(); ex.reserve(hh, ww, nb)
        st = torch.cuda.Stream(dev); ex.set_stream(st.cuda_stream)        # (a real stream: 0 would mean "the handle's own")
        kpx = torch.empty((nb, capx, 28), dtype=torch.uint8, device=dev); dex = torch.empty((nb, capx, 32), dtype=torch.uint8, device=dev)
        cnx = torch.zeros(nb, dtype=torch.int32, device=dev)
        run = lambda: ex.extract_batch_raw(vid.data_ptr(), nb, hh, ww, ww, hh * ww, True, kpx.data_ptr(), dex.data_ptr(), capx, cnx.data_ptr(), True)
        for _ in range(3):
            run()
        t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(dev); t0.record(st)
        for _ in range(10):
            run()
        t1.record(st); torch.cuda.synchronize(dev); ex.synchronize()
        out["frames_per_s_%s_device_resident" % name] = nb * 10 / (t0.elapsed_time(t1) * 1e-3)
        del ex, vid, kpx, dex, cnx
    # config 4: 2000 x 2000 descriptors of two frames extracted with the 2000-feature extractor, all on device
    fr = synth.video(2, 720, 1280, seed=9)
    e2 = os1_b200.ORBextractor(2000, 1.2, 8, 20, 7, device=dev.index)
    d0 = e2(fr[0])[1][:2000]; d1 = e2(fr[1])[1][:2000]
    a = torch.from_numpy(np.ascontiguousarray(d0)).to(dev); b = torch.from_numpy(np.ascontiguousarray(d1)).to(dev)
    o = [torch.zeros(len(d0), dtype=torch.int32, device=dev) for _ in range(3)]
    s = torch.cuda.current_stream(dev)
    call = lambda: matcher.hamming_knn2_raw(a.data_ptr(), len(d0), b.data_ptr(), len(d1), o[0].data_ptr(), o[1].data_ptr(), o[2].data_ptr(),
                                            device=dev.index, stream=s.cuda_stream)
    for _ in range(20):
        call()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    reps = 200
    torch.cuda.synchronize(dev); e0.record(s)
    for _ in range(reps):
        call()
    e1.record(s); torch.cuda.synchronize(dev)
    us = e0.elapsed_time(e1) * 1e3 / reps
    pairs = len(d0) * len(d1)
    out["hamming_knn2_%dx%d" % (len(d0), len(d1))] = {"us_per_call": us, "gpairs_per_s": pairs / us / 1e3, "gops_per_s": 24 * pairs / us / 1e3,
                                                       "note": "1 pair = 8 XOR + 8 POPC + 8 ADD on 32-bit words (SURVEY.md 8d); 2 launches per call"}
    call2 = lambda: matcher.hamming_knn2_noalloc_raw(a.data_ptr(), len(d0), b.data_ptr(), len(d1), o[0].data_ptr(), o[1].data_ptr(),
                                                     o[2].data_ptr(), device=dev.index, stream=s.cuda_stream)
    for _ in range(20):
        call2()
    torch.cuda.synchronize(dev); e0.record(s)
    for _ in range(reps):
        call2()
    e1.record(s); torch.cuda.synchronize(dev)
    us2 = e0.elapsed_time(e1) * 1e3 / reps
    out["hamming_knn2_noalloc_%dx%d" % (len(d0), len(d1))] = {"us_per_call": us2, "gpairs_per_s": pairs / us2 / 1e3, "gops_per_s": 24 * pairs / us2 / 1e3}
    # cross-frame sweep on one GPU: all pairs of 48 frames of a video, one launch
    from os1_b200 import shard
    F = 48
    vid = synth.video(F, 720, 1280, seed=77)
    e3 = os1_b200.ORBextractor(1500, 1.2, 8, 20, 7, device=dev.index)
    kp_, desc_, cnt_ = e3.extract_batch(vid)
    dsl = torch.from_numpy(desc_).to(dev); csl = torch.from_numpy(cnt_).to(dev)
    prs = shard.pair_partition(F, 0, 1)
    shard.sweep_pairs(dsl, csl, prs[:8], dev.index)
    torch.cuda.synchronize(dev); e0.record(s)
    acc = shard.sweep_pairs(dsl, csl, prs, dev.index)
    e1.record(s); torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1)
    dpairs = float((cnt_[prs[:, 0]].astype(np.float64) * cnt_[prs[:, 1]]).sum())
    out["hamming_sweep_%d_frames" % F] = {"frame_pairs": int(len(prs)), "descriptor_pairs": dpairs, "ms": ms, "gpairs_per_s": dpairs / ms / 1e6,
                                          "gops_per_s": 24 * dpairs / ms / 1e6, "accepted_matches": int(acc.sum().item())}
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
    dist_pairs = float((cnt.astype(np.float64) * (cnt - 1) / 2).sum())
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


# ------------------------------------------------------------------------------------------------------------------
def run_ours(args, rank, world, local_rank):
    import torch
    import os1_b200
    from os1_b200 import synth
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    B = args.batch
    frames_h = torch.from_numpy(synth.video(B, ROWS, COLS, seed=1000 + rank)).pin_memory()
    frames_d = frames_h.to(dev, non_blocking=False)                       # resident in HBM before the timed region
    ext = os1_b200.ORBextractor(*PARAMS, device=local_rank)
    cap = ext.max_keypoints()
    ext.reserve(ROWS, COLS, B)
    stream = torch.cuda.current_stream(dev)
    xstream = torch.cuda.Stream(dev)                 # the extractor's stream: kernels and timing events share it
    ext.set_stream(xstream.cuda_stream)
    kp_d = torch.empty((B, cap, 28), dtype=torch.uint8, device=dev)
    desc_d = torch.empty((B, cap, 32), dtype=torch.uint8, device=dev)
    cnt_d = torch.zeros(B, dtype=torch.int32, device=dev)
    kp_h = torch.empty((B, cap, 28), dtype=torch.uint8).pin_memory()
    desc_h = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
    cnt_h = torch.zeros(B, dtype=torch.int32).pin_memory()

    def step_device():
        ext.extract_batch_raw(frames_d.data_ptr(), B, ROWS, COLS, COLS, ROWS * COLS, True,
                              kp_d.data_ptr(), desc_d.data_ptr(), cap, cnt_d.data_ptr(), True)

    def step_e2e():
        ext.extract_batch_raw(frames_h.data_ptr(), B, ROWS, COLS, COLS, ROWS * COLS, False,
                              kp_h.data_ptr(), desc_h.data_ptr(), cap, cnt_h.data_ptr(), False)

    def barrier():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- device-resident throughput (`value`) + per-kernel durations for the roofline
    ext.set_timing(True)
    sampler = ClockSampler(local_rank); sampler.start()          # nvidia-smi needs ~0.3 s to come up: start before warm-up
    for _ in range(args.warmup):
        step_device()
    ext.synchronize()
    t_wait = time.time()
    while len(sampler.rows) < 2 and time.time() - t_wait < 3.0:    # keep the GPU busy until the sampler is reporting
        step_device(); ext.synchronize()
    pre_rows = list(sampler.rows[-2:])       # taken under load; used only if the timed region is shorter than a sampling period
    sampler.rows = []
    barrier()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    stage_ms = np.zeros(4)
    e0.record(xstream)
    for _ in range(args.steps):
        step_device()
        stage_ms += ext.stage_times_ms()        # waits for this step's last event: steps are serialised anyway
    e1.record(xstream)
    barrier()
    ms = e0.elapsed_time(e1)
    if not sampler.rows:
        sampler.rows = pre_rows
        clocks_src = "under load immediately before the timed region (it was shorter than the 50 ms sampling period)"
    else:
        clocks_src = "during the timed region"
    clocks = sampler.finish()
    clocks["sampled"] = clocks_src
    ext.synchronize()
    counts = cnt_d.cpu().numpy()
    if dist is not None:
        t = torch.tensor([ms], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t.item())
    ext.set_timing(False)

    # ---- end to end through the C ABI with pinned host buffers
    for _ in range(max(1, args.warmup // 2)):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    torch.cuda.synchronize(dev)
    e2e_s = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([e2e_s], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX); e2e_s = float(t.item())
    assert np.array_equal(cnt_h.numpy(), counts), "host and device paths disagree on keypoint counts"
    nkp = int(counts.sum())

    # ---- the path's one exchange step (N > 1): all-gather descriptor slabs over NCCL, then each rank matches its
    # share of the cross-frame pairs (BASELINE.json configs[4], scaled to the first frames of every rank)
    sweep = None
    if dist is not None:
        from os1_b200 import shard
        Fl = min(32, B)
        dsl = desc_d[:Fl].contiguous(); csl = cnt_d[:Fl].contiguous()
        shard.allgather_descriptors(dsl, csl)             # warm-up (NCCL channel set-up)
        barrier()
        s0 = torch.cuda.Event(enable_timing=True); s1 = torch.cuda.Event(enable_timing=True); s2 = torch.cuda.Event(enable_timing=True)
        s0.record(stream)
        d_all, c_all = shard.allgather_descriptors(dsl, csl)
        s1.record(stream)
        pairs = shard.pair_partition(Fl * world, rank, world)
        acc = shard.sweep_pairs(d_all, c_all, pairs, local_rank)
        s2.record(stream)
        barrier()
        tg = torch.tensor([s0.elapsed_time(s1), s1.elapsed_time(s2), float(len(pairs)), float(acc.sum().item())], device=dev, dtype=torch.float64)
        tmax = tg.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = tg.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        npairs = int(tsum[2].item())
        cmean = float(c_all.float().mean().item())
        sweep = {"frames": Fl * world, "frame_pairs": npairs, "allgather_ms": float(tmax[0].item()),
                 "allgather_bytes_per_rank": int(dsl.numel() + csl.numel() * 4),
                 "match_ms": float(tmax[1].item()), "accepted_matches": int(tsum[3].item()),
                 "gpairs_per_s": npairs * cmean * cmean / (float(tmax[1].item()) * 1e-3) / 1e9,
                 "gops_per_s": 24 * npairs * cmean * cmean / (float(tmax[1].item()) * 1e-3) / 1e9,
                 "note": "descriptor pairs/s over all ranks, one batched launch per rank incl. the acceptance test; NCCL all_gather_into_tensor"}

    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    fps = world * B * args.steps / (ms * 1e-3)
    e2e_fps = world * B * args.steps / e2e_s
    # roofline of the dominant kernel
    names = ["pyramid", "fast", "quadtree", "describe"]
    alg = algorithmic_bytes(ROWS, COLS, nkp / B)
    dom = int(np.argmax(stage_ms))
    per_launch_ms = stage_ms / args.steps
    peak, which = peaks()
    dom_bytes = alg[names[dom]] * B
    achieved = dom_bytes / (per_launch_ms[dom] * 1e-3) / 1e9
    whole = alg["total"] * B / (ms / args.steps * 1e-3) / 1e9
    # dram__bytes_read.sum + dram__bytes_write.sum per FRAME of this workload, from the ncu --set full capture of this
    # very command (profiles/r01f_ncu_full_raw.csv, 256 frames per launch); both scale with the frames per launch
    ncu_traffic_per_frame = {"pyramid": 720.72e6 / 256, "fast": 792.31e6 / 256, "quadtree": 271.65e6 / 256, "describe": 688.46e6 / 256}
    roof = {"bound": "hbm", "kernel": names[dom] + "_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": ncu_traffic_per_frame[names[dom]] * B,
            "traffic_source": "profiles/r01f_ncu_full_raw.csv (ncu --set full, same command, per launch)",
            "note": "FAST is bound by the integer ALU pipe and instruction issue (ncu: ALU 80 % busy, issue 73 %, DRAM 4 %), not by HBM: see DESIGN.md section 4", "peak_source": which + " (MEASURED_PEAKS.json hbm_gbs)" if which == "measured" else "fallback 6650 GB/s",
            "algorithmic_bytes_per_launch": dom_bytes,
            "kernel_ms_per_launch": {n: float(v) for n, v in zip(names, per_launch_ms)},
            "whole_step_achieved_gbs": whole, "whole_step_frac": whole / peak}
    line = {"metric": "ORB frames/sec", "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "frames_per_step_per_gpu": B, "keypoints_per_frame": nkp / B,
                       "l2": "inputs larger than L2 (%d MB of frames per step per GPU, plus %d MB of pyramid)" % (
                           B * ROWS * COLS >> 20, B * 1930000 >> 20),
                       "parallelism": "frame-sharded, %d rank(s), no data-path collective" % world},
            "e2e": {"value": e2e_fps, "unit": "frames/s", "h2d_bytes_per_step": int(B * ROWS * COLS),
                    "d2h_bytes_per_step": int(B * cap * 60 + B * 4 + 4)},
            "gpu_launches": int(args.steps * ext.launches_per_call()),
            "roofline": roof, "clocks": clocks}
    if sweep is not None:
        line["sweep"] = sweep
    if world == 1:
        try:
            line["extras"] = extras(torch, os1_b200, dev)
        except Exception as ex:
            line["extras"] = {"error": str(ex)}
        try:
            line["cpu_baseline"] = cpu_baseline_sample()
        except Exception as ex:           # the checker libraries are optional on the box; say so rather than fail the bench
            line["cpu_baseline"] = {"value": None, "unit": "frames/s", "cores": 0, "kind": "port", "sample": "unavailable: %s" % ex}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=256, help="frames per step per GPU")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
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
