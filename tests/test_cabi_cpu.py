"""CPU tests of the drop-in boundary: the product library loads without a GPU and exports every
symbol include/cugo_b200.h declares; without a device it refuses to create a solver (there is no
CPU fallback); the product package never imports the oracle."""
import ctypes as C
import os
import re

import pytest

from cugo_b200 import solver

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_symbols_exported():
    lib = solver.load_library()
    hdr = open(os.path.join(ROOT, "include", "cugo_b200.h")).read()
    declared = set(re.findall(r"\b(cugo_b200_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(solver.SYMBOLS)
    for name in declared:
        assert hasattr(lib, name), name


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(solver.CugoError):
        solver.BlockSolver(0)
    lib = solver.load_library()
    h = C.c_void_p()
    assert lib.cugo_b200_create(0, 0, C.byref(h)) == -2     # CUGO_B200_ENODEVICE
    assert lib.cugo_b200_create(0, 1, C.byref(h)) == -2     # the USE_FLOAT32 mode is a device mode too
    assert not h.value


def test_product_does_not_touch_oracle():
    pkg = os.path.join(ROOT, "cuda-bundle-adjustment_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")) or f == "Makefile":
                txt = open(os.path.join(d, f), errors="ignore").read()
                assert "ba_oracle" not in txt and "import oracle" not in txt and "oracle/" not in txt, os.path.join(d, f)


def test_cpp_facade_sample_builds_and_refuses_without_device(tmp_path):
    """include/cugo_b200.hpp (the reference-style C++ API) compiles with the host compiler alone, links
    only the C ABI library, and reports the missing device instead of falling back to a CPU path."""
    import subprocess
    import torch
    from cugo_b200 import graphio, synth
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "cuda-bundle-adjustment_b200"), "lib/sample_ba_from_bag"],
                          stdout=subprocess.DEVNULL)
    exe = os.path.join(ROOT, "cuda-bundle-adjustment_b200", "lib", "sample_ba_from_bag")
    assert os.path.exists(exe)
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "cuda-bundle-adjustment_b200"), "lib/sample_icp"],
                          stdout=subprocess.DEVNULL)
    exe_icp = os.path.join(ROOT, "cuda-bundle-adjustment_b200", "lib", "sample_icp")
    assert os.path.exists(exe_icp)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    bag = str(tmp_path / "tiny.bag")
    graphio.write_graph(synth.make_graph("tiny"), bag)
    r = subprocess.run([exe, bag, "3"], capture_output=True, text=True)
    assert r.returncode == 4 and "no CUDA device" in r.stderr
    # the pose-only sample (PlaneEdgeSet / LineEdgeSet, pose vertex set alone) likewise
    icp = str(tmp_path / "p.icp")
    graphio.write_icp(synth.make_icp_problem(n_poses=3, n_planes=40, n_lines=30), icp)
    r = subprocess.run([exe_icp, icp, "3"], capture_output=True, text=True)
    assert r.returncode == 4 and "no CUDA device" in r.stderr
