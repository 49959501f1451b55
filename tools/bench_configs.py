"""Secondary measurements on BASELINE.json's other configs (the headline bench.py line is config 1 batched):
wall-clock time of the reference-facing host calls on a B200, with the CPU oracle (port of the reference) timed beside
them on a bounded sample.  One JSON object per config on stdout.

    python tools/bench_configs.py [--no-cpu]
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from ydorbslam_b200 import OrbExtractor, OrbMatcher, synth  # noqa: E402

CPU = "--no-cpu" not in sys.argv
THREADS = os.cpu_count() or 1


def best_of(f, n=5, warm=2):
    for _ in range(warm):
        f()
    ts = []
    for _ in range(n):
        t0 = time.perf_counter()
        f()
        ts.append(time.perf_counter() - t0)
    return min(ts), sorted(ts)[len(ts) // 2]


def cpu_time(f, n=2):
    if not CPU:
        return None
    f()
    t0 = time.perf_counter()
    for _ in range(n):
        f()
    return (time.perf_counter() - t0) / n


def emit(**kw):
    print(json.dumps(kw), flush=True)


def config1():
    from oracle import orc

    img = synth.synth_frame(640, 480, 21)
    ex = OrbExtractor(1000, 1.2, 8, 20, 7, max_width=640, max_height=480, max_batch=1)
    best, med = best_of(lambda: ex.extractAndCompute(img), n=20, warm=5)
    p = orc.make_params(1000, 1.2, 8, 20)
    c = cpu_time(lambda: orc.extract(p, img), n=5)
    emit(config=1, workload="one 640x480 frame, 1000 kp, host call incl. H2D/D2H (latency-bound)", gpu_us_best=best * 1e6,
         gpu_us_median=med * 1e6, cpu_port_us_1thread=None if c is None else c * 1e6)
    ex.close()


def config2():
    from oracle import orc

    w, h, nf, bf = 1241, 376, 2000, 386.1448
    left, right = synth.synth_stereo_pair(w, h, 64)
    exl = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    exr = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    m = OrbMatcher()

    def gpu():
        kl, dl = exl.extractAndCompute(left)
        kr, dr = exr.extractAndCompute(right)
        return m.computeStereoMatches(exl, 0, exr, 0, kl, dl, kr, dr, bf, np.inf, 0)

    best, med = best_of(gpu, n=10, warm=3)
    # the same pair as ONE two-frame extract call on one handle (frames 0 and 1), then the stereo search
    ex2 = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=2)
    pair = np.stack([left, right])

    def gpu_pair():
        kps, desc, n = ex2.extract_batch(pair)
        return m.computeStereoMatches(ex2, 0, ex2, 1, kps[0, :n[0]], desc[0, :n[0]], kps[1, :n[1]], desc[1, :n[1]], bf, np.inf, 0)

    a, b = gpu(), gpu_pair()
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    best2, med2 = best_of(gpu_pair, n=10, warm=3)
    ex2.close()
    p = orc.make_params(nf, 1.2, 8, 20)

    def cpu():
        kl, dl, _ = orc.extract(p, left, cap=nf + 64)
        kr, dr, _ = orc.extract(p, right, cap=nf + 64)
        _, pl = orc.pyramid(p, left)
        _, pr = orc.pyramid(p, right)
        orc.stereo_match(p, w, h, pl, pr, kl, dl, kr, dr, bf, np.inf, 0)

    c = cpu_time(cpu, n=2)
    emit(config=2, workload="1241x376 stereo pair, 2000 kp/image, two extractAndCompute + computeStereoMatches (host calls)",
         gpu_ms_best=best * 1e3, gpu_ms_median=med * 1e3, pairs_per_s=1.0 / best, gpu_ms_best_one_call_pair=best2 * 1e3,
         gpu_ms_median_one_call_pair=med2 * 1e3, cpu_port_ms_1thread=None if c is None else c * 1e3)
    exl.close(); exr.close(); m.close()


def config3():
    import torch
    from oracle import orc

    B, w, h, nf = 64, 1920, 1080, 4000
    base = np.stack([synth.synth_frame(w, h, 300 + i) for i in range(4)])
    imgs = np.concatenate([base] * (B // 4))
    ex = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B)
    cap = ex.capacity
    dev = torch.device("cuda:0")
    d_img = torch.from_numpy(imgs).to(dev)
    d_kps = torch.zeros(B, cap, 28, dtype=torch.uint8, device=dev)
    d_desc = torch.zeros(B, cap, 32, dtype=torch.uint8, device=dev)
    d_n = torch.zeros(B, dtype=torch.int32, device=dev)
    st = torch.cuda.Stream()
    ex.set_stream(st.cuda_stream)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(st):
        for _ in range(2):
            ex.extract_batch_device(d_img.data_ptr(), w, h, B, d_kps.data_ptr(), d_desc.data_ptr(), cap, d_n.data_ptr())
        ex.set_profiling(True)
        e0.record(st)
        for _ in range(4):
            ex.extract_batch_device(d_img.data_ptr(), w, h, B, d_kps.data_ptr(), d_desc.data_ptr(), cap, d_n.data_ptr())
        e1.record(st)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 4
    stage = ex.stage_ms()
    ex.set_profiling(False)
    ex.set_stream(None)
    h_best, _ = best_of(lambda: ex.extract_batch(imgs), n=3, warm=1)
    # the same host call with the frames and the outputs in page-locked memory (what bench.py's e2e leg uses)
    from ydorbslam_b200._lib import KP_DTYPE, check, lib, ptr
    import ctypes as C

    p_img = torch.from_numpy(imgs).pin_memory()
    p_kps = torch.zeros(B * cap * KP_DTYPE.itemsize, dtype=torch.uint8).pin_memory()
    p_desc = torch.zeros(B * cap * 32, dtype=torch.uint8).pin_memory()
    p_n = torch.zeros(B, dtype=torch.int32).pin_memory()

    def pinned_call():
        check(lib().ydorb_extract_batch(ex._h, ptr(p_img.data_ptr()), w, h, w, C.c_size_t(h * w), B, ptr(p_kps.data_ptr()),
                                        ptr(p_desc.data_ptr()), cap, ptr(p_n.data_ptr())))

    hp_best, _ = best_of(pinned_call, n=3, warm=1)
    assert int(p_n.sum()) == int(d_n.sum().item())
    p = orc.make_params(nf, 1.2, 8, 20)
    c = cpu_time(lambda: orc.extract_batch(p, imgs[:THREADS], THREADS, cap=nf + 64), n=1)
    emit(config=3, workload=f"{B} frames 1920x1080, 4000 kp, one GPU (a 256-frame batch is 4 such launches per GPU at G=1)",
         device_frames_per_s=B / (ms * 1e-3), device_ms_per_batch=ms, stage_ms=dict(zip(["pyramid", "fast", "quadtree", "describe"], stage)),
         host_api_frames_per_s=B / h_best, host_api_pinned_frames_per_s=B / hp_best, keypoints_per_frame=float(d_n.float().mean().item()),
         cpu_port_frames_per_s=None if c is None else THREADS / c, cpu_threads=THREADS)
    ex.close()


def config5():
    from oracle import orc
    from test_gpu_search_stereo import _mps_array

    w, h, nf, nmp = 3840, 2160, 8000, 200_000
    img = synth.synth_frame(w, h, 500, n_rect=200)
    ex = OrbExtractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    kps, desc = ex.extractAndCompute(img)
    e_best, e_med = best_of(lambda: ex.extractAndCompute(img), n=10, warm=2)
    sf = ex.getScaleFactors()
    mp = synth.synth_map_points(kps, desc, nmp, w, h, 8, seed=3)
    mps = _mps_array(mp)
    rng = np.random.RandomState(5)
    right_x = np.where(rng.rand(nf) < 0.5, kps["x"] - rng.uniform(0, 30, nf), -1.0).astype(np.float32)
    occ = (rng.rand(nf) < 0.05).astype(np.uint8)
    m = OrbMatcher(0.8)
    fr = m.make_frame(kps, desc, right_x, bounds=(0, 0, w, h), scale_factors=sf)
    s_best, s_med = best_of(lambda: m.searchByProjectionInFrameAndMapPoint(fr, mps, mp["desc"], mp["obs_positive"], 3.0, kp_occupied=occ),
                            n=10, warm=2)
    p = orc.make_params(nf, 1.2, 8, 20)
    g = orc.make_geom(p, w, h)
    ce = cpu_time(lambda: orc.extract(p, img, cap=nf + 64), n=1)
    cs = cpu_time(lambda: orc.search_by_projection(g, kps, desc, right_x, occ, mps, mp["desc"], mp["obs_positive"], 3.0, 0.8), n=1)
    emit(config=5, workload="3840x2160, 8000 kp extractAndCompute, then searchByProjectionInFrameAndMapPoint vs 200k map points (host calls)",
         extract_ms_best=e_best * 1e3, extract_ms_median=e_med * 1e3, search_ms_best=s_best * 1e3, search_ms_median=s_med * 1e3,
         map_points_per_s=nmp / s_best, cpu_port_extract_ms_1thread=None if ce is None else ce * 1e3,
         cpu_port_search_ms_1thread=None if cs is None else cs * 1e3)
    fr.close(); m.close(); ex.close()


if __name__ == "__main__":
    for f in (config1, config2, config3, config5):
        f()
