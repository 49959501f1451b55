"""Seeded scenarios shared by the CPU (oracle self-check) and GPU (parity) matcher tests."""
import numpy as np

from ydorbslam_b200 import queries as Q
from ydorbslam_b200 import synth
from ydorbslam_b200._lib import KP_DTYPE

W, H = 640, 480
SF = np.array([np.float32(pow(np.float32(1.2), l)) for l in range(8)], np.float32)  # src/orbExtractor.cpp:329-332
SF2 = (SF * SF).astype(np.float32)
INV_SF2 = (np.float32(1.0) / SF2).astype(np.float32)


def synth_keypoints(n, seed, w=W, h=H):
    """Key points + descriptors without running the extractor (CPU tests have no GPU): uniform positions on a
    half-pixel lattice scaled per octave like extractAndCompute does (:391-396), random angles."""
    r = synth.rand_u64(seed, 21, n * 4)
    octave = (r[0::4] % np.uint64(8)).astype(np.int32)
    kps = np.zeros(n, KP_DTYPE)
    lx = (r[1::4] % np.uint64(w - 40)).astype(np.float32) + 20
    ly = (r[2::4] % np.uint64(h - 40)).astype(np.float32) + 20
    kps["x"] = np.floor(lx / SF[octave]) * SF[octave]
    kps["y"] = np.floor(ly / SF[octave]) * SF[octave]
    kps["octave"] = octave
    kps["size"] = (31 * SF[octave]).astype(np.int32)
    kps["angle"] = ((r[3::4] >> np.uint64(11)).astype(np.float64) / float(1 << 53) * 360.0).astype(np.float32)
    kps["response"] = 40
    kps["class_id"] = -1
    return kps, synth.synth_descriptors(n, seed + 5)


def window_scene(kps, desc, n_mp, seed, thd, dx_extra=8.0, w=W, h=H):
    """Projected map points + the per-point inputs the query builders take.  The reference's window predicate keeps a
    key point when |dx| > r and |dy| < r, so a point that should find its source key point is projected
    just OUTSIDE the radius in x (still inside the cell range) and inside it in y; its level is the source octave
    -1/0/+1 so that the different level filters of the members each keep some and drop some."""
    n = max(n_mp, 1)
    r = synth.rand_u64(seed, 31, n * 8)
    u = [(r[k::8] >> np.uint64(11)).astype(np.float64) / float(1 << 53) for k in range(8)]
    src = (r[0::8] % np.uint64(max(len(kps), 1))).astype(np.int64)
    near = (r[1::8] % np.uint64(5)) != 0
    level = np.clip(kps["octave"][src] + (r[2::8] % np.uint64(4)).astype(np.int64) - 2, 0, 7).astype(np.int32)
    level = np.where(near, level, (r[2::8] % np.uint64(8)).astype(np.int32)).astype(np.int32)
    rad = np.float32(thd) * SF[level]
    sign = np.where(r[3::8] & np.uint64(1), 1.0, -1.0)
    dx = sign * (rad + 0.05 + u[4] * dx_extra)
    dy = (u[5] - 0.5) * 2.2 * rad
    px = np.where(near, kps["x"][src] + dx, u[6] * w).astype(np.float32)
    py = np.where(near, kps["y"][src] + dy, u[7] * h).astype(np.float32)
    d = np.where(near[:, None], synth.flip_bits(desc[src], 0.08, seed + 1), synth.synth_descriptors(n, seed + 3))
    r2 = synth.rand_u64(seed, 32, n * 2)
    mp = dict(proj_x=px, proj_y=py, proj_right_x=(px - 10.0).astype(np.float32), level=level,
              active=((r2[0::2] % np.uint64(16)) != 0).astype(np.int32), obs_positive=((r2[1::2] % np.uint64(8)) != 0).astype(np.uint8),
              desc=np.ascontiguousarray(d), src=src)
    for k in list(mp):
        mp[k] = mp[k][:n_mp]
    mp["angle"] = synth.synth_source_angles(kps, mp["src"], seed + 17)
    rng = np.random.RandomState(seed)
    mp["right_x_kp"] = np.where(rng.rand(len(kps)) < 0.5, kps["x"] - 10.0 + rng.uniform(-1.5, 1.5, len(kps)) * dx_extra, -1.0).astype(np.float32)
    mp["occupied"] = (rng.rand(len(kps)) < 0.1).astype(np.uint8)
    return mp


def all_window_cases(kps, desc, n_mp, seed, thd=7.0):
    """(name, queries, params, kp_occupied, scene) for every M2/M3/M4 member."""
    mp = window_scene(kps, desc, n_mp, seed, thd)
    a = mp["active"].astype(bool)
    out = []
    for motion in ("forward", "backward", "none"):
        q, p = Q.last_and_current_frame(mp["proj_x"], mp["proj_y"], mp["proj_right_x"], mp["level"], mp["angle"], a,
                                        mp["obs_positive"].astype(bool), SF, thd, motion)
        out.append((f"last_{motion}", q, p, mp["occupied"], mp))
    q, p = Q.keyframe_and_current_frame(mp["proj_x"], mp["proj_y"], mp["level"], mp["angle"], a, SF, thd, 80)
    out.append(("keyframe", q, p, mp["occupied"], mp))
    q, p = Q.keyframe_and_current_frame(mp["proj_x"], mp["proj_y"], mp["level"], mp["angle"], a, SF, thd, 100, check_orientation=False)
    out.append(("keyframe_noori", q, p, None, mp))
    q, p = Q.in_sim(mp["proj_x"], mp["proj_y"], mp["level"], a, SF, int(thd))
    out.append(("in_sim", q, p, mp["occupied"], mp))
    q, p = Q.sim3_direction(mp["proj_x"], mp["proj_y"], mp["level"], a, SF, thd)
    out.append(("sim3_dir", q, p, None, mp))
    # fuseByProjection: the chi2 gates (:718-719) need |d| <= 2.45 sf while the window predicate needs |dx| > thd * sf
    # (SURVEY F9), so it can only ever accept with thd < 2.45; its scene uses thd = 1.2 and a tight x offset
    fp = window_scene(kps, desc, n_mp, seed + 1, 1.2, dx_extra=1.0)
    q, p = Q.fuse_by_projection(fp["proj_x"], fp["proj_y"], fp["proj_right_x"], fp["level"], fp["active"].astype(bool), SF, INV_SF2, 1.2)
    out.append(("fuse_proj", q, p, None, fp))
    q, p = Q.fuse_by_sim3(mp["proj_x"], mp["proj_y"], mp["level"], a, SF, thd)
    out.append(("fuse_sim3", q, p, None, mp))
    return out


def bow_scene(n1, n2, seed):
    """Two feature sets where ~60 % of side 2 are perturbed copies of side-1 features (same BoW node with high
    probability), with their FeatureVectors."""
    k1, d1 = synth_keypoints(n1, seed)
    k2, d2 = synth_keypoints(n2, seed + 1)
    r = synth.rand_u64(seed, 23, n2 * 2)
    src = (r[0::2] % np.uint64(max(n1, 1))).astype(np.int64)
    copy = (r[1::2] % np.uint64(5)) < 3
    if n1 and n2:
        d2 = np.where(copy[:, None], synth.flip_bits(d1[src], 0.04, seed + 2), d2)
        d2[:, 0] = np.where(copy, d1[src, 0], d2[:, 0])  # keep the node word of copies
        k2["angle"] = np.where(copy, np.mod(k1["angle"][src] + synth.synth_source_angles(k1, src, seed + 3) * 0.1, 360.0),
                               k2["angle"]).astype(np.float32)
    rng = np.random.RandomState(seed)
    v1 = (rng.rand(n1) < 0.8).astype(np.uint8)
    v2 = (rng.rand(n2) < 0.8).astype(np.uint8)
    fv1 = synth.synth_feature_vector(d1, seed)
    fv2 = synth.synth_feature_vector(d2, seed + 1)
    rx1 = np.where(rng.rand(n1) < 0.4, k1["x"] - 5, -1.0).astype(np.float32)
    rx2 = np.where(rng.rand(n2) < 0.4, k2["x"] - 5, -1.0).astype(np.float32)
    return dict(k1=k1, d1=np.ascontiguousarray(d1), v1=v1, fv1=fv1, rx1=rx1, k2=k2, d2=np.ascontiguousarray(d2), v2=v2,
                fv2=fv2, rx2=rx2)


def fundamental(seed):
    """A plausible F (row-major float32) and epipole for searchForTriangulation gates."""
    rng = np.random.RandomState(seed)
    # scaled so that num^2 / den^2 of :814 straddles 3.841 * sf2 (the reference squares the denominator)
    F = np.array([[0.2, -0.4, 80.0], [0.4, 0.2, -140.0], [60.0, 100.0, 4000.0]], np.float32) * (1 + 0.2 * rng.rand(3, 3).astype(np.float32))
    return F.astype(np.float32), (np.float32(300.0 + 50 * rng.rand()), np.float32(200.0 + 50 * rng.rand()))


def init_scene(n_clusters, n_noise, seed, area_mode, window=30, w=W, h=H):
    """Two frames for the searchForInitialization extra.  Frame 1 has clusters of 1-4 nearby key points whose
    descriptors are noisy copies of one base descriptor; frame 2 has one key point per cluster carrying the base
    descriptor, placed where the window predicate of `area_mode` keeps it for every member (mode 0 = the reference's
    getKeyPointsInArea: |dx| > r and |dy| < r; mode 1 = ORB-SLAM2's |dx| < r and |dy| < r), so members
    compete for it: later members with a smaller distance steal it, others are blocked by the recorded distance.
    Unrelated level-0 key points in frame 2 supply second-best distances; ~30 % of frame 1 sits above level 0."""
    rng = np.random.RandomState(seed)
    base = synth.synth_descriptors(n_clusters, seed + 1)
    k1, d1, k2, d2 = [], [], [], []
    for c in range(n_clusters):
        cx, cy = rng.uniform(60, w - 60), rng.uniform(60, h - 60)
        ang = rng.uniform(0, 360)
        members = rng.randint(1, 5)
        for _ in range(members):
            octv = 0 if rng.rand() < 0.7 else rng.randint(1, 8)
            k1.append((np.floor(cx + rng.uniform(-1.5, 1.5)), np.floor(cy + rng.uniform(-1.5, 1.5)), octv,
                       (ang + rng.uniform(-3, 3)) % 360.0))
            d1.append(synth.flip_bits(base[c : c + 1], rng.choice([0.02, 0.05, 0.08, 0.12]), seed * 1000 + len(k1))[0])
        if area_mode == 0:
            dx = rng.choice([-1.0, 1.0]) * (window + 2.0 + rng.uniform(0, 6))
        else:
            dx = rng.uniform(-10, 10)
        rot = 0.0 if rng.rand() < 0.8 else rng.uniform(0, 360)  # most pairs share one rotation, some are outliers
        k2.append((np.floor(cx + dx), np.floor(cy + rng.uniform(-10, 10)), 0, (ang + 20.0 + rot) % 360.0))
        d2.append(base[c])
    noise = synth.synth_descriptors(max(n_noise, 1), seed + 2)
    for i in range(n_noise):
        k2.append((np.floor(rng.uniform(20, w - 20)), np.floor(rng.uniform(20, h - 20)), 0 if rng.rand() < 0.8 else 1, rng.uniform(0, 360)))
        d2.append(noise[i])

    def pack(ks):
        out = np.zeros(len(ks), KP_DTYPE)
        for i, (x, y, o, a) in enumerate(ks):
            out[i]["x"], out[i]["y"], out[i]["octave"], out[i]["angle"] = x, y, o, a
        out["size"] = 31
        out["response"] = 40
        out["class_id"] = -1
        return out

    kps1, kps2 = pack(k1), pack(k2)
    perm = rng.permutation(len(kps2))  # frame 2 in arbitrary order
    return kps1, np.ascontiguousarray(np.array(d1, np.uint8)), kps2[perm], np.ascontiguousarray(np.array(d2, np.uint8)[perm])


CAM = dict(fx=np.float32(718.856), fy=np.float32(718.856), cx=np.float32(320.0), cy=np.float32(240.0), bf=np.float32(386.1448),
           baseline=np.float32(386.1448) / np.float32(718.856))


def last_frame_scene(kps, desc, n_last, seed, thd):
    """A 'last frame' for searchByProjectionInLastAndCurrentFrame against a current frame at the identity pose: world
    positions whose projections (computed in float32 exactly as src/orbMatcher.cpp:86-91 does) land where window_scene
    wants them, octaves / angles of the last frame's key points, outlier flags, missing map points (NaN)."""
    mp = window_scene(kps, desc, n_last, seed, thd)
    rng = np.random.RandomState(seed + 100)
    z = rng.uniform(2.0, 20.0, n_last).astype(np.float32)
    z[rng.rand(n_last) < 0.03] *= -1  # behind the camera: :92 drops them
    X = ((mp["proj_x"].astype(np.float64) - float(CAM["cx"])) / float(CAM["fx"]) * z).astype(np.float32)
    Y = ((mp["proj_y"].astype(np.float64) - float(CAM["cy"])) / float(CAM["fy"]) * z).astype(np.float32)
    pos = np.stack([X, Y, z], 1).astype(np.float32)
    has_mp = rng.rand(n_last) > 0.08
    outlier = (rng.rand(n_last) < 0.05).astype(np.uint8)
    # the reference's own arithmetic, one float32 rounding per operation
    px = (CAM["fx"] * X / z + CAM["cx"]).astype(np.float32)
    py = (CAM["fy"] * Y / z + CAM["cy"]).astype(np.float32)
    prx = (px - CAM["bf"] / z).astype(np.float32)
    in_image = (px >= 0) & (px < W) & (py >= 0) & (py < H)
    active = has_mp & (outlier == 0) & (z >= 0) & in_image
    last_kps = np.zeros(n_last, KP_DTYPE)
    last_kps["octave"] = mp["level"]
    last_kps["angle"] = mp["angle"]
    last_kps["size"] = 31
    last_kps["class_id"] = -1
    pos_nan = pos.copy()
    pos_nan[~has_mp, 0] = np.nan
    return dict(last_kps=last_kps, pos=pos_nan, proj_x=px, proj_y=py, proj_right_x=prx, active=active, outlier=outlier,
                obs_positive=mp["obs_positive"], desc=mp["desc"], right_x_kp=mp["right_x_kp"], occupied=mp["occupied"])
