"""GPU: the C++ drop-in drivers (2012-04_stanford_cme213_b200/host) -- the reference programs' command lines,
printed lines and files -- and, where it was built, the reference's own fp binary run beside them."""
import os
import re
import subprocess

import numpy as np
import pytest

import oracle
from oracle import generators as g
from conftest import ROOT, golden

pytestmark = pytest.mark.gpu
BIN = os.path.join(ROOT, "2012-04_stanford_cme213_b200", "bin")


def write_inputs(d, a, s, k, x, iters):
    with open(os.path.join(d, "a.txt"), "w") as f:
        f.write("%d %d %d %d\n" % (len(a), len(s), len(x), iters))
        f.write(" ".join(repr(float(v)) for v in a) + "\n")
        f.write(" ".join(str(int(v)) for v in s) + "\n")
        f.write(" ".join(str(int(v)) for v in k) + "\n")
    with open(os.path.join(d, "x.txt"), "w") as f:
        f.write(" ".join(repr(float(v)) for v in x) + "\n")


def test_fp_driver_fixture_and_checker(tmp_path):
    f = golden("fp_small.npz")
    write_inputs(tmp_path, f["a"], f["s"], f["k"], f["x"], 1)
    r = subprocess.run([os.path.join(BIN, "fp"), "a.txt", "x.txt", "check"], cwd=tmp_path, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    assert re.search(r"The running time of my code for 1 iterations is: [0-9.e+-]+ milliseconds\.", r.stdout)  # Final.pdf p.3
    assert "Elements beyond tolerance 0.01: 0" in r.stdout
    b = np.array(open(tmp_path / "b.txt").read().split(), dtype=np.float64)
    assert np.allclose(b, f["b"], atol=2e-5, rtol=0)  # text output carries 6 significant digits
    r = subprocess.run([os.path.join(BIN, "check_output"), "a.txt", "x.txt", "b.txt"], cwd=tmp_path, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    m = re.search(r"relative l2/linf\s+(\S+)\s+(\S+)\s+(\S+)\s+(\S+)", r.stdout)
    assert m and float(m.group(3)) < 1e-5 and float(m.group(4)) < 1e-5


def test_fp_driver_larger_input_vs_reference_binary(tmp_path):
    a, s, k, x = g.fp_inputs(200000, 3000, 5000, seed=5, dtype=np.float32)
    write_inputs(tmp_path, a, s, k, x, 4)
    r = subprocess.run([os.path.join(BIN, "fp"), "a.txt", "x.txt"], cwd=tmp_path, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    b = np.array(open(tmp_path / "b.txt").read().split(), dtype=np.float64)
    want = oracle.sparse_mv_scan(a.astype(np.float64), s, k, x.astype(np.float64), 4)
    norms = oracle.error_norms(want, b)
    assert norms[2] < 2e-5, norms  # 6 printed digits + fp32
    # bad input is reported, not asserted
    open(tmp_path / "bad.txt", "w").write("4 3 3 1\n1 1 1 1\n0 5 4\n0 1 2 1\n")
    r = subprocess.run([os.path.join(BIN, "fp"), "bad.txt", "x.txt"], cwd=tmp_path, capture_output=True, text=True, timeout=60)
    assert r.returncode == 1 and "s" in r.stderr
    ref = os.path.join(ROOT, "oracle", "_ref", "fp_ref")
    if os.path.exists(ref):
        # the reference program itself on this GPU.  Its warp-synchronous kernel has no defined behaviour on
        # sm_70+, so this is informational: it must run, and we record how far it is off.
        os.rename(tmp_path / "b.txt", tmp_path / "b_ours.txt")
        rr = subprocess.run([ref, "a.txt", "x.txt"], cwd=tmp_path, capture_output=True, text=True, timeout=120)
        if rr.returncode == 0 and os.path.exists(tmp_path / "b.txt"):
            b_ref = np.array(open(tmp_path / "b.txt").read().split(), dtype=np.float64)
            print("reference fp binary on this GPU: rel-L2 vs double oracle =", oracle.error_norms(want, b_ref)[2])


def test_pagerank_driver(tmp_path):
    r = subprocess.run([os.path.join(BIN, "pagerank"), "300002", "8", "20", "7"], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "gpu graph propagate took" in r.stdout and "Worked! CUDA and reference output match." in r.stdout


def test_radixsort_driver(tmp_path):
    # hw/hw4/programming/radixsort.cpp:162-214 with the GPU sort in the place of radixSortParallel
    r = subprocess.run([os.path.join(BIN, "radixsort"), "3000001", "8"], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "n_elements: 3000001" in r.stdout and "parallel radix:" in r.stdout
    assert "Worked! GPU radix sort and std::sort output match." in r.stdout
    r = subprocess.run([os.path.join(BIN, "radixsort"), "100000", "4"], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "Worked!" in r.stdout, r.stdout + r.stderr
