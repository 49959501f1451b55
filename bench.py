#!/usr/bin/env python3
"""bench.py -- ORB extract+match frames/s (640x480, 1000 kp) and Hamming Gmatch/s on B200.

    python bench.py --gpus N --steps K --warmup W                 # this repo's CUDA path (libydorb.so)
    python bench.py --impl reference --gpus N --steps K --warmup W  # reference CPU path (oracle port) on host cores
    torchrun --nproc-per-node N ... bench.py --gpus N ...          # one rank per GPU, frames sharded, no collective

One step = one pass of the hot path over one batch of synthetic frames per GPU:
  OrbExtractor::extractAndCompute(1000, 1.2, 8, 20, 7) on B frames of 640x480, then the Hamming top-2 + ratio test
  (rule of src/orbMatcher.cpp:319-336) of every frame's descriptors against the previous frame's.
Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for the byte counts behind `roofline`.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H, NFEAT = 640, 480, 1000
EXTRACTOR_ARGS = (NFEAT, 1.2, 8, 20, 7)
RATIO, TH_LOW = 0.6, 50
HAMMING_NQ, HAMMING_NDB = 2000, 1_000_000
E2E_CALL_FRAMES = int(os.environ.get("YDORB_E2E_CALL_FRAMES", "128"))  # frames per host call of the threaded e2e leg
E2E_THREADS = int(os.environ.get("YDORB_E2E_THREADS", "4"))  # host threads (one extractor + matcher handle each) of the e2e leg


def level_sizes(w, h, n_levels=8, scale=1.2):
    out = []
    for l in range(n_levels):
        inv = np.float32(pow(float(np.float32(scale)), -l))
        out.append((int(np.rint(np.float32(w) * inv)), int(np.rint(np.float32(h) * inv))))
    return out


def algorithmic_bytes(w, h):
    """per-stage algorithmic bytes of one frame (1 B/px)."""
    px = [a * b for a, b in level_sizes(w, h)]
    pyr = sum(px[:-1]) + sum(px[1:])   # resize reads + writes
    fast = sum(px)                     # FAST reads every level once
    blur = 2 * sum(px)                 # blur read + write (the reference blurs whole levels)
    return {"pyramid": pyr, "fast": fast, "quadtree": 0, "describe": blur, "total": pyr + fast + blur}


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.splitlines():
            f = [t.strip() for t in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def make_frames(batch, rank):
    from ydorbslam_b200 import synth

    # One textured canvas per rank; frame i is a window sliding 2 px per frame (and wobbling vertically), so
    # consecutive frames are distinct images that overlap -> the frame-to-frame matcher finds real matches.
    # Dense texture: ~10k FAST candidates at level 0, every level overshoots its quota (truncation + tie paths run).
    canvas = synth.synth_frame(W + 2 * batch, H + 8, 100000 * rank + 17, n_rect=64 + batch // 2, contrast=2)
    return np.stack([canvas[(i * 3) % 8 : (i * 3) % 8 + H, 2 * i : 2 * i + W] for i in range(batch)])


# ------------------------------------------------------------------------------------------------ reference arm
def cpu_kind():
    """("reference", note) when oracle/_ref (the reference's own translation units compiled unmodified against the OpenCV
    stand-in, oracle/refshim/Makefile) is present, else ("port", note)."""
    try:
        from oracle import ref

        if ref.available():
            ref.lib()
            return "reference", ("oracle/_ref: /root/reference/src/orbExtractor.cpp compiled unmodified (-std=c++11 -O3) against the "
                                 "OpenCV stand-in (scalar resize / FAST / GaussianBlur models, no SIMD OpenCV in this image); "
                                 "dense top-2 + ratio test = the rule of src/orbMatcher.cpp:319-336 in oracle/ydorb_oracle.cpp")
    except Exception:  # noqa: BLE001
        pass
    return "port", "oracle/ydorb_oracle.cpp (oracle/_ref not shipped)"


def cpu_extract_match(frames, threads, kind="port"):
    """The reference CPU path on `frames` with `threads` host threads: a fresh OrbExtractor(1000,1.2,8,20,7) per frame
    (kind "reference": the compiled reference; "port": the oracle restatement), then top-2 + ratio test of every frame
    against the previous one.  Returns seconds."""
    from oracle import orc

    t0 = time.perf_counter()
    if kind == "reference":
        from oracle import ref

        kps, desc, n = ref.extract_batch(frames, threads, *EXTRACTOR_ARGS, cap=NFEAT + 64)
    else:
        p = orc.make_params(*EXTRACTOR_ARGS[:4])
        kps, desc, n = orc.extract_batch(p, frames, threads, cap=NFEAT + 64)
    for b in range(len(frames)):
        a = (b - 1) % len(frames)
        t = orc.top2(desc[b, : n[b]], desc[a, : n[a]], 0, threads)
        orc.ratio_accept(t, RATIO, TH_LOW)
    return time.perf_counter() - t0


def cv2_primitives(frames, threads):
    """Secondary bracket: only the OpenCV calls the reference makes per frame -- resize x7,
    copyMakeBorder x8, FAST, GaussianBlur x8 -- through real SIMD OpenCV (cv2), in the reference's call order
    (oracle/cv2_flow.py).  No quadtree, angles or descriptors, so `frames_per_s` is an UPPER bound on what a build of the
    reference against real OpenCV reaches per `threads`.  FAST is timed twice: per cell from Python as the reference
    calls it (~800 calls per frame, dominated by call overhead here) and once per level, scaled by the (36/30)^2 pixels
    the overlapping cells re-read -- the bound uses the latter.  Returns a dict or None without cv2."""
    try:
        import cv2

        from oracle import cv2_flow
    except Exception:  # noqa: BLE001
        return None
    cv2.setNumThreads(threads)
    inv = cv2_flow.ctor(*EXTRACTOR_ARGS[:3])[1]
    det = cv2.FastFeatureDetector_create(EXTRACTOR_ARGS[3], True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    t_pyr = t_cell = t_level = t_blur = 0.0
    pad = cv2_flow.PAD
    for img in frames:
        t0 = time.perf_counter()
        levels = cv2_flow.pyramid(img, inv, EXTRACTOR_ARGS[2])
        t1 = time.perf_counter()
        for pl in levels:
            cv2_flow.level_candidates(pl[pad:-pad, pad:-pad], EXTRACTOR_ARGS[3], EXTRACTOR_ARGS[3])
        t2 = time.perf_counter()
        for pl in levels:
            det.detect(np.ascontiguousarray(pl[pad - 3:-(pad - 3), pad - 3:-(pad - 3)]))
        t3 = time.perf_counter()
        for pl in levels:
            cv2.GaussianBlur(pl, (7, 7), 2, None, 2, cv2.BORDER_REFLECT_101)
        t4 = time.perf_counter()
        t_pyr += t1 - t0
        t_cell += t2 - t1
        t_level += t3 - t2
        t_blur += t4 - t3
    n = len(frames)
    per_frame = (t_pyr + 1.44 * t_level + t_blur) / n
    return {"threads": threads, "frames_per_s": 1.0 / per_frame, "ms_per_frame": {
        "pyramid": 1e3 * t_pyr / n, "fast_per_cell_calls": 1e3 * t_cell / n, "fast_whole_levels": 1e3 * t_level / n,
        "gaussian_blur": 1e3 * t_blur / n}}


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return 0
    threads = os.cpu_count() or 1
    kind, note = cpu_kind()
    frames = make_frames(args.batch, 0)
    tp = cpu_extract_match(frames[: min(threads, args.batch)], threads, kind)  # page in / warm up, and a speed probe
    per_frame = tp / min(threads, args.batch)
    # the GPU arm's own batch per step when the whole run then stays within a few minutes, else a bounded sample
    budget = 150.0
    sample = args.batch if per_frame * args.batch * (args.steps + args.warmup) <= budget else \
        max(threads, int(budget / ((args.steps + args.warmup) * per_frame)))
    frames = frames[:sample]
    for _ in range(max(args.warmup - 1, 0)):
        cpu_extract_match(frames, threads, kind)
    times = [cpu_extract_match(frames, threads, kind) for _ in range(args.steps)]
    total = sum(times)
    value = sample * args.steps / total
    line = {
        "impl": "reference", "metric": "ORB extract+match frames/s (640x480, 1000 kp)", "value": value, "unit": "frames/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": f"{sample} synthetic 640x480 frames/step"
                               + ("" if sample == args.batch else " (bounded sample of the GPU arm's batch)")
                               + ": extractAndCompute(1000,1.2,8,20,7) + Hamming top-2 ratio match vs previous frame, CPU",
                   "parallelism": f"{threads} host threads", "same_batch_as_gpu_arm": sample == args.batch},
        "cpu_baseline": {"value": value, "unit": "frames/s", "cores": threads, "kind": kind,
                         "sample": f"{sample} frames x {args.steps} steps; {note}"},
        "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------------ GPU arm
def run_gpu(args):
    import torch
    import torch.distributed as dist

    from ydorbslam_b200 import OrbExtractor, OrbMatcher, synth
    from ydorbslam_b200._lib import lib

    rank, world, local = dist_env()
    # stdout carries exactly one JSON line: anything libraries print there (NCCL's version banner under NCCL_DEBUG) goes
    # to stderr instead while the run lasts
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    B, K, Wm = args.batch, args.steps, args.warmup

    frames = make_frames(B, rank)
    ex = OrbExtractor(*EXTRACTOR_ARGS, max_width=W, max_height=H, max_batch=B, device=local)
    m = OrbMatcher(RATIO, device=local)
    cap = ex.capacity
    stream = torch.cuda.Stream(device=dev)
    ex.set_stream(stream.cuda_stream)
    m.set_stream(stream.cuda_stream)

    d_img = torch.from_numpy(frames).to(dev)
    d_kps = torch.zeros(B, cap, 28, dtype=torch.uint8, device=dev)
    d_desc = torch.zeros(B, cap, 32, dtype=torch.uint8, device=dev)
    d_n = torch.zeros(B, dtype=torch.int32, device=dev)
    d_qf = torch.arange(B, dtype=torch.int32, device=dev)
    d_df = (torch.arange(B, dtype=torch.int32, device=dev) - 1) % B
    d_top2 = torch.zeros(B, cap, 4, dtype=torch.int32, device=dev)
    d_acc = torch.zeros(B, cap, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()

    # Steps are issued back to back.  The matcher has its own stream and the step's results alternate between two
    # buffer sets, so the Hamming match of step i (ALU-bound) runs beside the extraction of step i + 1 (latency-bound):
    # extraction i waits for match i - 2 (same buffers), match i for extraction i.  YDORB_BENCH_PIPE=0: one stream.
    pipe = os.environ.get("YDORB_BENCH_PIPE", "1") != "0"
    mstream = torch.cuda.Stream(device=dev)
    sets = [(d_kps, d_desc, d_n, d_top2, d_acc)]
    if pipe:
        sets.append(tuple(torch.zeros_like(x) for x in sets[0]))
        m.set_stream(mstream.cuda_stream)
    ev_ext = [torch.cuda.Event() for _ in sets]
    ev_match = [torch.cuda.Event() for _ in sets]
    counter = [0]

    def step():
        i = counter[0] % len(sets)
        counter[0] += 1
        kps_, desc_, n_, top2_, acc_ = sets[i]
        if pipe:
            stream.wait_event(ev_match[i])
        ex.extract_batch_device(d_img.data_ptr(), W, H, B, kps_.data_ptr(), desc_.data_ptr(), cap, n_.data_ptr())
        if pipe:
            ev_ext[i].record(stream)
            mstream.wait_event(ev_ext[i])
        m.top2_frames_device(desc_.data_ptr(), n_.data_ptr(), cap, d_qf.data_ptr(), d_df.data_ptr(), B, top2_.data_ptr(),
                             acc_.data_ptr(), RATIO, TH_LOW)
        if pipe:
            ev_match[i].record(mstream)

    def join():  # the timed stream waits for the matches still in flight
        if pipe:
            for e in ev_match:
                stream.wait_event(e)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.cuda.stream(stream):
        for _ in range(Wm):
            step()
    ex.sync()
    barrier()
    launches0 = lib().ydorb_kernel_launches()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        e0.record(stream)
        for _ in range(K):
            step()
        join()
        e1.record(stream)
    barrier()
    ms_total = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    launches = lib().ydorb_kernel_launches() - launches0
    # sustained figure: the same step looped for >= 2 s of device time (the 20-step region above lasts ~45 ms, a burst)
    sustained = None
    if args.sustain_s > 0:
        ks = max(K, int(args.sustain_s / max(ms_total / K * 1e-3, 1e-6)) + 1)
        s_sampler = ClockSampler(local)
        if rank == 0:
            s_sampler.start()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            s0.record(stream)
            for _ in range(ks):
                step()
            join()
            s1.record(stream)
        barrier()
        ts_ms = torch.tensor([s0.elapsed_time(s1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ts_ms, op=dist.ReduceOp.MAX)
        sustained = {"value": world * B * ks / (float(ts_ms.item()) * 1e-3), "unit": "frames/s", "steps": ks,
                     "seconds": float(ts_ms.item()) * 1e-3, "clocks": s_sampler.stop() if rank == 0 else None}
    # per-stage times: the same K steps again with the handle's stage events on (serial stages, one stream); the timed
    # region above runs the production path, which overlaps two sub-batches on two streams
    ex.set_profiling(True)
    barrier()
    pipe = False  # the serial pass: matcher back on the extractor's stream, one buffer set
    m.set_stream(stream.cuda_stream)
    sets = sets[:1]
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        p0.record(stream)
        for _ in range(K):
            step()
        p1.record(stream)
    barrier()
    serial_ms = p0.elapsed_time(p1) / K
    stage_ms = ex.stage_ms()
    ex.set_profiling(False)
    ex.sync()  # raises if a device-side status flag (capacity overflow) was set
    n_kp = d_n.cpu().numpy()
    n_acc = int(d_acc.sum().item())
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    value = world * B * K / (ms_total * 1e-3)

    # ---- e2e: host buffers through the host-pointer C-ABI calls, H2D/D2H inside the timed region.
    # E2E_THREADS host threads each own an extractor + matcher handle and a contiguous share of every step's frames and
    # run their K steps back to back (the reference drives its left and right extractor from two concurrent threads,
    # src/frame.cpp:84-87): one handle's upload overlaps another's kernels.  The single-thread figure is reported too.
    from ydorbslam_b200._lib import KP_DTYPE, TOP2_DTYPE, check, ptr
    import ctypes as C
    import threading

    ex.set_stream(None)
    m.set_stream(None)
    h_img = torch.from_numpy(frames).pin_memory()
    h_kps = torch.zeros(B, cap, 28, dtype=torch.uint8).pin_memory()
    h_desc = torch.zeros(B, cap, 32, dtype=torch.uint8).pin_memory()
    h_n = torch.zeros(B, dtype=torch.int32).pin_memory()
    h_top2 = torch.zeros(B, cap, 4, dtype=torch.int32).pin_memory()
    h_acc = torch.zeros(B, cap, dtype=torch.uint8).pin_memory()

    def make_worker(exh, mh, f0, n, outs, stride=1, first=0):
        """Frames [f0, f0 + n) of steps first, first + stride, ... through the two synchronous host calls."""
        qf = np.arange(n, dtype=np.int32)
        df = ((np.arange(n) - 1) % n).astype(np.int32)
        o_kps, o_desc, o_n, o_top2, o_acc = outs

        def run(steps):
            for _ in range(first, steps, stride):
                check(lib().ydorb_extract_batch(exh._h, C.c_void_p(h_img[f0:].data_ptr()), W, H, W, C.c_size_t(W * H), n,
                                                C.c_void_p(o_kps[f0:].data_ptr()), C.c_void_p(o_desc[f0:].data_ptr()), cap,
                                                C.c_void_p(o_n[f0:].data_ptr())))
                check(lib().ydorb_match_frames(mh._h, exh._h, ptr(qf), ptr(df), n, C.c_float(RATIO), TH_LOW,
                                               C.c_void_p(o_top2[f0:].data_ptr()), C.c_void_p(o_acc[f0:].data_ptr())))

        return run

    def timed(workers, steps):
        errs = []

        def guard(w):
            try:
                w(steps)
            except Exception as e:  # noqa: BLE001
                errs.append(e)

        th = [threading.Thread(target=guard, args=(w,)) for w in workers]
        t0 = time.perf_counter()
        for x in th:
            x.start()
        for x in th:
            x.join()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if errs:
            raise errs[0]
        return dt

    ke = max(min(K, 10), 1)

    def e2e_measure(workers, warm):
        timed(workers, warm)
        barrier()
        t = torch.tensor([timed(workers, ke)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return world * B * ke / float(t.item())

    outs0 = (h_kps, h_desc, h_n, h_top2, h_acc)
    e2e_single = e2e_measure([make_worker(ex, m, 0, B, outs0)], max(Wm, 1))
    assert np.array_equal(h_n.numpy(), n_kp), "host-API and device-API keypoint counts differ"
    # Threaded: every host call carries E2E_CALL_FRAMES frames (128: below that the kernels of a call are too small to
    # fill the GPU -- tools/e2e_batch_probe.py: 4 threads x 64 frames 110.8 k frames/s, 4 x 128 127.3 k), so a step is
    # B / 128 calls, each on its own thread and handle, and T / (B / 128) steps are in flight at a time, each with its
    # own pinned result buffers.  The synchronous entry points poll while they wait: keep the ranks' threads within
    # the host's cores.
    T = max(1, min(E2E_THREADS, B // 8, max(1, (os.cpu_count() or 4) // world)))
    call = min(B, max(8, E2E_CALL_FRAMES))
    slots = max(1, min(T, B // call))          # calls (threads) per step
    pipes = max(1, T // slots)                 # steps in flight
    T = slots * pipes
    bounds = [B * t // slots for t in range(slots + 1)]
    outs = [outs0] + [tuple(torch.zeros_like(x).pin_memory() for x in outs0) for _ in range(pipes - 1)]
    handles = [(OrbExtractor(*EXTRACTOR_ARGS, max_width=W, max_height=H, max_batch=bounds[t % slots + 1] - bounds[t % slots],
                             device=local), OrbMatcher(RATIO, device=local)) for t in range(T)]
    for o in outs:
        o[2].zero_()
    workers = [make_worker(handles[t][0], handles[t][1], bounds[t % slots], bounds[t % slots + 1] - bounds[t % slots],
                           outs[t // slots], pipes, t // slots) for t in range(T)]
    e2e_value = e2e_measure(workers, max(Wm, pipes))
    for o in outs[: min(pipes, ke)]:
        assert np.array_equal(o[2].numpy(), n_kp), "threaded host-API and device-API keypoint counts differ"
    del handles, workers, outs
    h2d = B * W * H + 2 * 4 * B
    d2h = B * cap * (28 + 32) + 4 * B + B * cap * (16 + 1)

    # ---- secondary metric: Hamming top-2, 2000 queries x 1M database rows sharded over the ranks, through the
    # library's own multi-GPU path (csrc/shard.cu: ydorb_comm_create + ydorb_hamming_top2_sharded_device): ONE kernel per
    # GPU and call -- shard scan, peer-memory push of the 16 B/query records over NVLink, merge + ratio test
    ham = None
    if not args.skip_hamming:
        from ydorbslam_b200 import sharding

        rows = HAMMING_NDB // world
        base = rank * rows
        shards_np = [synth.synth_descriptors(rows, 1000 + r) for r in range(world)]
        q_np, _ = synth.synth_query_descriptors(np.concatenate(shards_np), HAMMING_NQ, 7)
        d_db = torch.from_numpy(shards_np[rank]).to(dev)
        d_q = torch.from_numpy(q_np).to(dev)
        d_out = torch.zeros(HAMMING_NQ, 4, dtype=torch.int32, device=dev)
        d_ok = torch.zeros(HAMMING_NQ, dtype=torch.uint8, device=dev)
        m.set_stream(stream.cuda_stream)
        uid = [sharding.comm_unique_id() if (rank == 0 and world > 1) else None]
        if world > 1:
            dist.broadcast_object_list(uid, src=0)  # host-side plumbing only: ships the 128-byte ncclUniqueId
        comm = sharding.Communicator(m, rank, world, HAMMING_NQ, uid[0])

        def ham_step():
            sharding.sharded_top2_device(comm, d_q.data_ptr(), HAMMING_NQ, d_db.data_ptr(), rows, base, RATIO, TH_LOW,
                                         d_out.data_ptr(), d_ok.data_ptr())

        # correctness before timing, on EVERY rank: the merged result must equal one scan of the whole database on this
        # GPU (ydorb_hamming_top2_device over all shards + ratio test), field by field
        with torch.cuda.stream(stream):
            ham_step()
            ham_step()  # second call: the other parity of the double-buffered exchange
        comm.check()
        d_full = torch.from_numpy(np.concatenate(shards_np)).to(dev)
        d_ref = torch.zeros(HAMMING_NQ, 4, dtype=torch.int32, device=dev)
        d_ref_ok = torch.zeros(HAMMING_NQ, dtype=torch.uint8, device=dev)
        m.top2_device(d_q.data_ptr(), HAMMING_NQ, d_full.data_ptr(), rows * world, 0, d_ref.data_ptr())
        m.ratio_accept_device(d_ref.data_ptr(), HAMMING_NQ, d_ref_ok.data_ptr(), RATIO, TH_LOW)
        m.sync()
        got, want = d_out.cpu().numpy(), d_ref.cpu().numpy()
        assert np.array_equal(got[:, :3], want[:, :3]), f"rank {rank}: sharded top-2 differs from the single scan"
        assert np.array_equal(d_ok.cpu().numpy(), d_ref_ok.cpu().numpy()), f"rank {rank}: ratio-test flags differ"
        del d_full
        ham_true = int((got[: HAMMING_NQ // 2, 0] >= 0).sum())

        h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 20
        with torch.cuda.stream(stream):
            for _ in range(3):
                ham_step()
            barrier()
            h0.record(stream)
            for _ in range(reps):
                ham_step()
            h1.record(stream)
        barrier()
        comm.check()
        t = torch.tensor([h0.elapsed_time(h1) / reps], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ham_ms = float(t.item())
        ham = {"metric": "Hamming top-2 Gmatch/s", "value": HAMMING_NQ * rows * world / (ham_ms * 1e-3) / 1e9, "unit": "Gpairs/s",
               "ms": ham_ms, "workload": f"{HAMMING_NQ} queries x {rows * world} database descriptors, {world} shard(s)",
               "exchange": comm.exchange, "kernels_per_call": 1 if comm.exchange != "nccl" else 3,
               "verified": "merged result == single scan of the whole database on every rank (indices, distances, ratio flags)",
               "scaling": "strong", "int_ops_per_pair": {"lop3": 16, "vimnmx": 2.5, "popc": 4, "imad": 4},
               "accepted": int(d_ok.sum().item()), "queries_with_match": ham_true}
        comm.close()
        m.set_stream(None)

    # ---- integer-pipe peaks measured live on this device (tools/int_peaks.py, csrc/diag.cu): roofline denominators
    # of the kernels that are issue-bound rather than HBM-bound (DESIGN.md "Integer roofline")
    int_peaks = None
    if rank == 0:
        import ctypes as C2

        rates = (C2.c_double * 5)()
        check(lib().ydorb_diag_int_peaks(local, rates, 5))
        int_peaks = dict(zip(["popc", "lop3", "vimnmx3_s16x2", "imad", "lop3_imad_mix"], [r / 1e9 for r in rates]))

    # ---- CPU baseline beside it (rank 0, N=1 only): the reference's CPU path on the host cores, bounded sample of ~12 s,
    # plus the two figures BASELINE.md 3 asks for: reference-faithful threading (ONE thread per extractor,
    # src/frame.cpp:84-87) and the share real SIMD OpenCV would take (cv2 primitives only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        kind, note = cpu_kind()
        probe = frames[: min(threads, B)]
        tp = cpu_extract_match(probe, threads, kind)
        per_frame = tp / len(probe)
        want = max(len(probe), int(12.0 / max(per_frame, 1e-6)))
        done, ts = 0, 0.0
        while done < want and ts < 30.0:
            chunk = frames[: min(B, want - done)]
            ts += cpu_extract_match(chunk, threads, kind)
            done += len(chunk)
        n1 = max(4, min(B, int(3.0 / max(per_frame * threads, 1e-6))))
        t1 = cpu_extract_match(frames[:n1], 1, kind)
        cv_one = cv2_primitives(frames[:16], 1)
        cv_all = cv2_primitives(frames[:32], threads)
        cpu = {"value": done / ts, "unit": "frames/s", "cores": threads, "kind": kind,
               "sample": f"{done} frames (the batch's frames, repeated) extract + match vs previous frame, {threads} threads, "
                         f"{ts:.1f} s; {note}",
               "one_thread_per_extractor": {"value": n1 / t1, "unit": "frames/s", "cores": 1,
                                            "sample": f"{n1} frames, one extractor on one thread (the reference's own threading, "
                                                      "src/frame.cpp:84-87)"},
               "cv2_primitives_only": None if cv_one is None else {
                   "one_thread": cv_one, "all_threads": cv_all,
                   "what": "UPPER bound for a real-OpenCV build: only cv2 4.13 resize / copyMakeBorder / FAST / GaussianBlur in "
                           "the reference's call order (oracle/cv2_flow.py), no quadtree / angle / descriptor"}}

    if rank == 0:
        peak, peak_kind = measured_peak()
        ab = algorithmic_bytes(W, H)
        names = ["pyramid", "fast", "quadtree", "describe"]
        match_ms = serial_ms - sum(stage_ms)
        dom = int(np.argmax(stage_ms))
        dom_bytes = ab[names[dom]] * B
        achieved = dom_bytes / (stage_ms[dom] * 1e-3) / 1e9 if stage_ms[dom] > 0 else 0.0
        step_gbs = ab["total"] * B / (ms_total / K * 1e-3) / 1e9
        line = {
            "metric": "ORB extract+match frames/s (640x480, 1000 kp)", "value": value, "unit": "frames/s", "n_gpus": world,
            "steps": K, "warmup": Wm, "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": f"{B} synthetic 640x480 frames/step/GPU: extractAndCompute(1000,1.2,8,20,7) + Hamming top-2 "
                                   "ratio match vs previous frame (1000x1000 per frame)",
                       "parallelism": f"frames sharded over {world} GPU(s), no collective",
                       "l2": f"working set {B} x (input 0.31 MB + pyramid 1.2 MB) > 126 MB L2, no flush"},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "timing": f"host wall clock around the synchronous host-pointer C-ABI calls (ydorb_extract_batch + "
                              f"ydorb_match_frames), pinned buffers; {T} host threads, each with its own extractor + matcher "
                              f"handle: a step is {slots} call(s) of {B // slots} frames on {slots} thread(s), {pipes} step(s) in "
                              "flight, each thread's calls back to back",
                    "host_threads": T, "frames_per_call": B // slots, "steps_in_flight": pipes,
                    "single_thread_value": e2e_single},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": names[dom], "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": None, "peak_kind": peak_kind,
                         "algorithmic_bytes_per_launch": dom_bytes, "launch_ms": stage_ms[dom]},
            "stage_ms": dict(zip(names, stage_ms), match=match_ms, serial_step=serial_ms,
                             note="stages serialised on one stream with events (second pass of K steps); the timed step overlaps two sub-batches"),
            "roofline_step": {"algorithmic_bytes_per_frame": ab["total"], "achieved": step_gbs, "peak": peak, "unit": "GB/s",
                              "frac": step_gbs / peak},
            "keypoints_per_frame": float(n_kp.mean()), "accepted_matches_per_frame": n_acc / B,
            "hamming": ham, "cpu_baseline": cpu, "int_peaks_gops": int_peaks,
        }
        # Integer roofline of the Hamming kernels (DESIGN.md "Integer roofline").  The kernel computes the 256-bit
        # distance with a carry-save adder tree: per descriptor pair 16 LOP3 + 3 VIMNMX on the ALU pipe (the binding
        # one) and 4 POPC on the XU pipe -- not SURVEY 8(d)'s naive 8 XOR + 8 POPC + 8 IADD, whose POPC-pipe bound
        # (popc_peak / 8 pairs/s) this restatement exceeds; `vs_naive_popc_bound` says by how much.
        def ham_roofline(gpairs, ngpu):
            alu_peak = int_peaks["lop3"] * ngpu
            return {"bound": "alu pipe (LOP3/VIMNMX)", "ops_per_pair": {"alu": 18.5, "popc": 4, "imad": 4},
                    "achieved": gpairs * 18.5, "peak": alu_peak, "unit": "G ALU thread-ops/s", "frac": gpairs * 18.5 / alu_peak,
                    "popc_pipe_frac": gpairs * 4 / (int_peaks["popc"] * ngpu),
                    "vs_naive_popc_bound": gpairs * 8 / (int_peaks["popc"] * ngpu)}

        if ham and int_peaks:
            ham["int_roofline"] = ham_roofline(ham["value"], world)
        if int_peaks:
            pairs = float(n_kp.astype(np.float64).dot(np.roll(n_kp, 1).astype(np.float64)))  # frame b vs frame b-1
            line["int_roofline"] = dict(kernel="k_top2_frames", **ham_roofline(pairs / (match_ms * 1e-3) / 1e9, 1))
        line["sustained"] = sustained
        # every stage against the measured HBM peak (algorithmic bytes of SURVEY 8d), and the fused lower bound SURVEY 8d
        # also names: input + outputs only, 367 200 B/frame
        tr = {}
        traffic_path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
        if os.path.exists(traffic_path):
            tr = json.load(open(traffic_path))
        sr = {}
        for i, nm in enumerate(names):
            gbs = ab[nm] * B / (stage_ms[i] * 1e-3) / 1e9 if stage_ms[i] > 0 else 0.0
            sr[nm] = {"ms": stage_ms[i], "algorithmic_bytes_per_frame": ab[nm], "achieved_gbs": gbs, "frac_hbm": gbs / peak,
                      "dram_bytes_per_frame_ncu": tr.get("bytes_per_frame", {}).get(nm)}
        mbytes = 2 * 32 * NFEAT + 17 * NFEAT  # both descriptor sets read once, top-2 record + flag written
        sr["match"] = {"ms": match_ms, "algorithmic_bytes_per_frame": mbytes,
                       "achieved_gbs": mbytes * B / (match_ms * 1e-3) / 1e9, "frac_hbm": mbytes * B / (match_ms * 1e-3) / 1e9 / peak}
        line["stage_roofline"] = sr
        fused = W * H + NFEAT * (28 + 32)
        line["roofline_step"]["fused_lower_bound_bytes_per_frame"] = fused
        line["roofline_step"]["fused_lower_bound_ms_per_step"] = fused * B / (peak * 1e9) * 1e3
        line["roofline_step"]["frac_of_fused_bound"] = (fused * B / (peak * 1e9) * 1e3) / (ms_total / K)
        # FAST is integer-issue bound, not HBM bound: warp instructions per frame from the committed ncu capture of the
        # same launch shape x frames/s of the stage, against the issue rate (4 schedulers x 1 warp-instruction per clock
        # per SM) at the SM clock sampled during the timed region; the ALU-pipe utilisation (the binding pipe: VIMNMX / LOP3 /
        # IADD3 issue at half rate) is the figure ncu measured for the same launch
        wi = tr.get("warp_inst_per_frame", {}).get("fast")
        if wi and clocks and clocks.get("sm_mhz"):
            sms = torch.cuda.get_device_properties(dev).multi_processor_count
            f_hz = clocks["sm_mhz"] * 1e6
            rate = wi * B / (stage_ms[1] * 1e-3)
            alu = tr.get("alu_pipe_pct_ncu", {}).get("fast")
            px = sum(a * b for a, b in level_sizes(W, H))
            line["fast_int_roofline"] = {
                "kernel": "k_fast_blocks", "bound": "instruction issue / ALU pipe", "warp_inst_per_frame": wi,
                "thread_inst_per_level_pixel": wi * 32 / px, "issue_rate_gwarp_inst_s": rate / 1e9,
                "issue_peak_gwarp_inst_s": sms * 4 * f_hz / 1e9, "frac_issue": rate / (sms * 4 * f_hz),
                "frac_alu_pipe_ncu": None if not alu else alu / 100.0,
                "algorithmic_ops": "per tested pixel pair: reject about 15 packed ALU ops (vertical + horizontal antipodal pairs: 2 shifts, "
                                   "6 min/max, 7 add / compare) on every pair; exact score 106 (two 9-arc trees 2 x 32, two 16-way reductions 2 x 8, 10 shifts, "
                                   "16 compare / select) on the pairs that pass (DESIGN.md 4)",
                "source": tr.get("source")}
        # The whole step against the same ceiling: warp instructions of all five stages (same capture) per step / the
        # timed step, over the issue rate of the GPU.  This, not HBM, is the roofline the path as built sits under.
        wis = tr.get("warp_inst_per_frame", {})
        if all(k in wis for k in names + ["match"]) and clocks and clocks.get("sm_mhz"):
            sms = torch.cuda.get_device_properties(dev).multi_processor_count
            peak_issue = sms * 4 * clocks["sm_mhz"] * 1e6
            total = sum(wis[k] for k in names + ["match"])
            line["issue_roofline_step"] = {
                "bound": "instruction issue (4 warp-instructions per clock per SM)", "warp_inst_per_frame": total,
                "by_stage": {k: wis[k] for k in names + ["match"]},
                "achieved_gwarp_inst_s": total * B / (ms_total / K * 1e-3) / 1e9, "peak_gwarp_inst_s": peak_issue / 1e9,
                "frac": total * B / (ms_total / K * 1e-3) / peak_issue,
                "ms_per_step_at_full_issue": total * B / peak_issue * 1e3, "source": tr.get("source")}
        if os.path.exists(traffic_path):
            # dram__bytes_read.sum + dram__bytes_write.sum per frame of each stage kernel, from the committed
            # `ncu --set full` capture (profiles/): scaled to this launch's batch
            if names[dom] in tr.get("bytes_per_frame", {}):
                line["roofline"]["traffic"] = tr["bytes_per_frame"][names[dom]] * B
                line["roofline"]["traffic_source"] = tr.get("source")
        real_stdout.write(json.dumps(line) + "\n")
        real_stdout.flush()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--skip-hamming", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sustain-s", type=float, default=2.0, help="also loop the step for this many seconds (0 = off)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
