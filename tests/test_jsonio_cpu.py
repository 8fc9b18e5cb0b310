"""Row f2: the reference's JSON graph schema (samples/sample_ba_from_file/main.cpp:78-160) read and written without
OpenCV, by the Python module (cugo_b200/jsonio.py) and by the C++ header (include/cugo_b200_json.hpp), and the
comparison report of samples/sample_comparison_with_g2o/main.cpp:101-148 (cugo_b200/compare.py)."""
import json
import os
import subprocess

import numpy as np
import pytest

from cugo_b200 import compare, graphio, jsonio, pack, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FIELDS = ("pose_q", "pose_t", "pose_fixed", "lm_xyz", "lm_fixed", "mono_p", "mono_l", "mono_meas", "mono_info",
          "stereo_p", "stereo_l", "stereo_meas", "stereo_info", "camera")


def _same(a, b):
    for f in FIELDS:
        np.testing.assert_array_equal(getattr(a, f), getattr(b, f), err_msg=f)


def test_json_round_trip_is_bit_exact(tmp_path):
    g = synth.make_graph("small")
    p = str(tmp_path / "g.json")
    jsonio.write_json(g, p)
    doc = json.load(open(p))                       # a standard JSON document with the reference's keys
    assert set(doc) == {"fx", "fy", "cx", "cy", "bf", "pose_vertices", "landmark_vertices", "monocular_edges", "stereo_edges"}
    assert set(doc["pose_vertices"][0]) == {"id", "fixed", "q", "t"}
    assert set(doc["stereo_edges"][0]) == {"vertexP", "vertexL", "measurement", "information"}
    g2, pid, lid = jsonio.read_json(p)
    _same(g, g2)
    np.testing.assert_array_equal(pid, np.arange(g.n_poses))
    # BAG1 <-> JSON: the binary format of the harness carries the same fields
    bag = str(tmp_path / "g.bag")
    graphio.write_graph(g2, bag)
    _same(g, graphio.read_graph(bag))


def test_json_arbitrary_ids_are_mapped_in_ascending_order(tmp_path):
    """ORB-SLAM dumps carry keyframe / map-point ids, not 0..n-1: positions follow ascending id, the order
    VertexSet::generateEstimateData walks its std::map (src/optimisable_graph.hpp:84-126)."""
    g = synth.make_graph("tiny")
    rng = np.random.default_rng(3)
    pid = np.sort(rng.choice(10000, g.n_poses, replace=False))
    lid = np.sort(rng.choice(1000000, g.n_landmarks, replace=False)) + 20000
    p = str(tmp_path / "ids.json")
    jsonio.write_json(g, p, pose_ids=pid, landmark_ids=lid)
    # shuffle the records of the file: order in the file must not matter
    doc = json.load(open(p))
    for k in ("pose_vertices", "landmark_vertices", "monocular_edges", "stereo_edges"):
        perm = rng.permutation(len(doc[k]))
        doc[k] = [doc[k][i] for i in perm]
        if k == "monocular_edges":
            mono_perm = perm
        if k == "stereo_edges":
            stereo_perm = perm
    json.dump(doc, open(p, "w"))
    g2, pid2, lid2 = jsonio.read_json(p)
    np.testing.assert_array_equal(pid2, pid)
    np.testing.assert_array_equal(lid2, lid)
    np.testing.assert_array_equal(g2.pose_q, g.pose_q)
    np.testing.assert_array_equal(g2.lm_xyz, g.lm_xyz)
    np.testing.assert_array_equal(g2.mono_p, g.mono_p[mono_perm])
    np.testing.assert_array_equal(g2.stereo_meas, g.stereo_meas[stereo_perm])
    # the packed problem (what the engine sees) has the same vertices; edges are a permutation
    a, b = pack.pack(g), pack.pack(g2)
    np.testing.assert_array_equal(a.poses, b.poses)
    assert a.n_edges == b.n_edges
    with pytest.raises(ValueError):
        doc["monocular_edges"][0]["vertexP"] = 123456789
        json.dump(doc, open(p, "w"))
        jsonio.read_json(p)


def test_cpp_reader_agrees_with_python(tmp_path):
    """include/cugo_b200_json.hpp parses what jsonio writes (and what json.dump writes: other whitespace, exponents,
    extra keys) and writes what jsonio reads back bit-exactly.  --convert needs no GPU."""
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "cuda-bundle-adjustment_b200"), "lib/sample_ba_from_file"],
                          stdout=subprocess.DEVNULL)
    exe = os.path.join(ROOT, "cuda-bundle-adjustment_b200", "lib", "sample_ba_from_file")
    g = synth.make_graph("small")
    src, dst = str(tmp_path / "a.json"), str(tmp_path / "b.json")
    jsonio.write_json(g, src)
    doc = json.load(open(src))
    doc["comment"] = {"nested": [1, 2, {"x": "y\\\"z"}], "flag": True}       # unknown keys are skipped
    doc["pose_vertices"][0]["extra"] = [1e-3, -2.5E+2]
    json.dump(doc, open(src, "w"), indent=1)
    subprocess.check_call([exe, src, "--convert", dst], stdout=subprocess.DEVNULL)
    g2, _, _ = jsonio.read_json(dst)
    _same(g, g2)
    bad = str(tmp_path / "bad.json")
    open(bad, "w").write('{"fx": 1, "fy": 1, "cx": 0, "cy": 0, "bf": 0, "pose_vertices": [{"id": 1, "q": [0,0,0,1}]}')
    r = subprocess.run([exe, bad, "--convert", dst], capture_output=True, text=True)
    assert r.returncode == 4 and "cugo json" in r.stderr


def test_comparison_report():
    rng = np.random.default_rng(0)
    pa, la = rng.standard_normal((5, 7)), rng.standard_normal((9, 3))
    pb, lb = pa.copy(), la.copy()
    pb[:, :4] += 1e-3; pb[:, 4:] -= 2e-3; lb += 3e-3
    r = compare.rmse_report(pa, la, pb, lb)
    assert r["rotation"] == pytest.approx(2e-3) and r["translation"] == pytest.approx(2e-3 * np.sqrt(3))
    assert r["landmark"] == pytest.approx(3e-3 * np.sqrt(3))
    txt, _ = compare.report([10.0, 9.0, 8.5], (pa, la), [10.0, 9.0], (pb, lb), "CPU", "GPU")
    lines = txt.splitlines()
    assert lines[0].split("|")[0].strip() == "iteration" and "N/A" in lines[3] and "Rotation" in txt
