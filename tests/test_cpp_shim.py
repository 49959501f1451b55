"""The C++ host shim (include/ydorb.hpp): on CPU it must compile and link against libydorb.so; on a B200 the program
tests/cpp/shim_check.cpp drives it like the reference's stereo Frame constructor and every output is compared with the
CPU oracle (bit-exact)."""
import pathlib
import subprocess

import numpy as np
import pytest

ROOT = pathlib.Path(__file__).resolve().parents[1]


def _build(tmp):
    from ydorbslam_b200 import build

    so = build.build()
    exe = tmp / "shim_check"
    subprocess.check_call(["g++", "-std=c++11", "-O2", "-Wall", "-Wextra", "-Werror", f"-I{ROOT / 'include'}",
                           str(ROOT / "tests" / "cpp" / "shim_check.cpp"), "-o", str(exe), f"-L{so.parent}", "-lydorb",
                           f"-Wl,-rpath,{so.parent}"])
    return exe


def test_shim_compiles_and_links(tmp_path):
    exe = _build(tmp_path)
    # without inputs the program stops at its first file read (exit code 2): proves it loads libydorb.so and starts
    assert subprocess.run([str(exe), str(tmp_path), str(tmp_path)], capture_output=True).returncode == 2


@pytest.mark.gpu
def test_shim_matches_oracle(tmp_path):
    from oracle import orc
    from tests import bow_cases
    from ydorbslam_b200 import synth
    from ydorbslam_b200._lib import KP_DTYPE, TOP2_DTYPE
    from ydorbslam_b200.vocabulary import write_dbow3_binary

    exe = _build(tmp_path)
    w, h, nf, levels_up = 640, 480, 1000, 2
    bf, baseline = np.float32(40.0), np.float32(0.08)
    left, right = synth.synth_stereo_pair(w, h, 77)
    ind, outd = tmp_path / "in", tmp_path / "out"
    ind.mkdir()
    outd.mkdir()
    np.array([w, h, nf, levels_up], np.int32).tofile(ind / "meta.bin")
    np.array([bf, baseline], np.float32).tofile(ind / "cam.bin")
    left.tofile(ind / "left.bin")
    right.tofile(ind / "right.bin")
    r = synth.rand_u64(5, 1, 60 * 3).astype(np.float64) / 2.0**64
    q = np.zeros((60, 5), np.float32)
    q[:, 0], q[:, 1], q[:, 2] = r[0::3] * w, r[1::3] * h, 5 + r[2::3] * 60
    q[:, 3] = np.tile([-1, 0, 2, 3, 1, 5], 10)
    q[:, 4] = np.tile([-1, 7, 4, 3, -1, 2], 10)
    q.tofile(ind / "area_queries.bin")
    v = bow_cases.tree(10, 3, 0.0, False)
    write_dbow3_binary(ind / "voc.dbow3", v)

    res = subprocess.run([str(exe), str(ind), str(outd)], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    assert "shim ok" in res.stdout

    p = orc.make_params(nf, 1.2, 8, 20)
    okl, odl, _ = orc.extract(p, left)
    okr, odr, _ = orc.extract(p, right)
    kl = np.fromfile(outd / "kps_left.bin", KP_DTYPE)
    kr = np.fromfile(outd / "kps_right.bin", KP_DTYPE)
    dl = np.fromfile(outd / "desc_left.bin", np.uint8).reshape(-1, 32)
    dr = np.fromfile(outd / "desc_right.bin", np.uint8).reshape(-1, 32)
    assert np.array_equal(kl, okl) and np.array_equal(kr, okr)
    assert np.array_equal(dl, odl) and np.array_equal(dr, odr)
    assert np.array_equal(np.fromfile(outd / "scale_factors.bin", np.float32), np.array(p.sf[:8], np.float32))
    levels, pl = orc.pyramid(p, left)
    _, pr = orc.pyramid(p, right)
    pw, ph, nl = np.fromfile(outd / "pyr3_dims.bin", np.int32)
    assert nl == 8 and np.array_equal(np.fromfile(outd / "pyr3.bin", np.uint8).reshape(ph, pw), levels[3][19:-19, 19:-19])
    # computeStereoMatches: maxD = bf / baseline
    orx, odp = orc.stereo_match(p, w, h, pl, pr, okl, odl, okr, odr, bf, np.float32(bf / baseline), 0)
    rx = np.fromfile(outd / "right_x.bin", np.float32)
    assert np.array_equal(rx.view(np.uint32), orx.view(np.uint32))
    assert np.array_equal(np.fromfile(outd / "depth.bin", np.float32).view(np.uint32), odp.view(np.uint32))
    assert (rx >= 0).sum() > 50
    # getKeyPointsInArea
    g = orc.make_geom(p, w, h)
    area = np.fromfile(outd / "area.bin", np.int32)
    pos = 0
    total = 0
    for x, y, rad, mn, mx in q:
        n = area[pos]
        ref = orc.keypoints_in_area(g, okl, x, y, rad, int(mn), int(mx))
        assert n == len(ref) and np.array_equal(area[pos + 1 : pos + 1 + n], ref)
        pos += 1 + n
        total += n
    assert pos == len(area) and total > 0
    # top-2 + ratio rule, single distance
    ref = orc.top2(odl, odr)
    t2 = np.fromfile(outd / "top2.bin", TOP2_DTYPE)
    assert np.array_equal(t2, ref)
    assert np.array_equal(np.fromfile(outd / "accept.bin", np.uint8), orc.ratio_accept(ref, 0.6, 50))
    assert np.fromfile(outd / "dist0.bin", np.int32)[0] == orc.hamming(odl[0], odr[0])
    # computeBoW through the file-loaded vocabulary
    ov = orc.Vocabulary.from_arrays(v["k"], v["L"], 0, 0, v["child_start"], v["children"], v["descriptors"], v["weights"], v["word_ids"])
    b = orc.voc_transform(ov, odl, levels_up)
    assert np.array_equal(np.fromfile(outd / "bow_ids.bin", np.int32), b["bow_ids"])
    assert np.array_equal(np.fromfile(outd / "bow_values.bin", np.float64).view(np.uint64), b["bow_values"].view(np.uint64))
    assert np.array_equal(np.fromfile(outd / "fv_nodes.bin", np.int32), b["fv_node_ids"])
    assert np.array_equal(np.fromfile(outd / "fv_start.bin", np.int32), b["fv_node_start"])
    assert np.array_equal(np.fromfile(outd / "fv_items.bin", np.int32), b["fv_items"])
