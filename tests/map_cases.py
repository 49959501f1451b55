"""Seeded inputs for the map-point rows, shared by the CPU and GPU tests."""
import numpy as np

from ydorbslam_b200 import mappoints, synth


def observation_sets(n_points, seed, max_obs=70):
    """CSR descriptor sets: every point is a random 256-bit prototype seen n times with a few bits flipped; some points
    have 0/1/2 observations, some contain exact duplicates (ties between medians), one is large (> 32 rows)."""
    rng = np.random.RandomState(seed)
    counts = rng.randint(0, 12, n_points)
    counts[rng.rand(n_points) < 0.1] = 0
    counts[rng.rand(n_points) < 0.05] = rng.randint(20, max_obs)
    if n_points > 3:
        counts[1], counts[2], counts[3] = 1, 2, max_obs
    start = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
    desc = np.zeros((start[-1], 32), np.uint8)
    for p in range(n_points):
        n = counts[p]
        if n == 0:
            continue
        proto = rng.randint(0, 256, 32).astype(np.uint8)
        flips = (rng.rand(n, 256) < rng.choice([0.02, 0.08, 0.3])).astype(np.uint8)
        rows = np.packbits(np.unpackbits(proto)[None, :] ^ flips, axis=1)
        if n >= 4 and rng.rand() < 0.4:
            rows[rng.randint(n)] = rows[rng.randint(n)]  # duplicate observation
        desc[start[p] : start[p + 1]] = rows
    return desc, start


def camera_and_points(n, seed, w=1241, h=376, n_levels=8, scale=1.2):
    """A KITTI-like camera with a random small rotation, and map points scattered so that every gate of
    isInCameraFrustum (behind the camera, outside the image, too near / too far, oblique normal) fails for some."""
    rng = np.random.RandomState(seed)
    ang = rng.uniform(-0.2, 0.2, 3)
    cx, sx, cy_, sy, cz, sz = np.cos(ang[0]), np.sin(ang[0]), np.cos(ang[1]), np.sin(ang[1]), np.cos(ang[2]), np.sin(ang[2])
    R = (np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]]) @ np.array([[cy_, 0, sy], [0, 1, 0], [-sy, 0, cy_]])
         @ np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])).astype(np.float32)
    t = rng.uniform(-2, 2, 3).astype(np.float32)
    cam = mappoints.Camera.make(R, t, 718.856, 718.856, 607.1928, 185.2157, (0.0, 0.0, float(w), float(h)), 386.1448, scale, n_levels)
    origin = np.array(cam.origin[:], np.float32)
    mp = np.zeros(n, mappoints.MAPPOINT3D_DTYPE)
    # points in camera coordinates, mostly in front, then moved to world: Pw = R^T (Pc - t)
    pc = np.stack([rng.uniform(-40, 40, n), rng.uniform(-12, 12, n), rng.uniform(-5, 60, n)], 1)
    pw = (R.T.astype(np.float64) @ (pc - t.astype(np.float64)).T).T
    mp["pos"] = pw.astype(np.float32)
    view = mp["pos"] - origin
    dist = np.linalg.norm(view, axis=1) + 1e-6
    nrm = view / dist[:, None] + rng.normal(0, 0.5, (n, 3))
    nrm /= np.linalg.norm(nrm, axis=1)[:, None] + 1e-9
    mp["normal"] = nrm.astype(np.float32)
    mp["max_distance"] = (dist * rng.uniform(0.5, 3.0, n)).astype(np.float32)
    mp["min_distance"] = (mp["max_distance"] / np.float32(scale ** (n_levels - 1))).astype(np.float32)
    return cam, mp
