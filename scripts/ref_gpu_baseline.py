"""The reference's OWN GPU programs on this B200, beside the drop-in drivers.

oracle/_ref/fp_ref and oracle/_ref/pagerank_ref are hw/hw_final/programming/fp.cu and hw/hw1/programming/pagerank.cu
compiled unmodified for sm_100a (oracle/Makefile).  This script runs them and bin/fp / bin/pagerank on the same inputs:
  * PageRank: the homework's own size (N = 2^21, avg_edges 8, 20 iterations; the reference seeds with time(NULL), so its
    graph differs in its random columns but not in shape);
  * final project: the 15 benchmark shapes (n, p, q, N) of hw/hw_final/programming/Final_1_Dong-Bang_Tsai.pdf pp.1-3 with
    seeded inputs in the recipe of aux/readMM.py:31-60 (the SuiteSparse matrices themselves are not in the checkout).
Both fp programs print "The running time of my code for N iterations is: X milliseconds."; both results are compared
with the instructor's double-precision recurrence (reference_spMVscan-released.cu:38-54 as restated in oracle.c).
Usage (GPU box): python scripts/ref_gpu_baseline.py [--only cant,pwtk] > profiles/r02_reference_gpu_programs.md"""
import argparse
import os
import re
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
BIN = os.path.join(ROOT, "2012-04_stanford_cme213_b200", "bin")
REF = os.path.join(ROOT, "oracle", "_ref")

# name: (n, p, q, N) -- Final_1_Dong-Bang_Tsai.pdf pp.1-3; N from Final_Report_DongBang_Tsai.tex:237-251
SHAPES = {
    "cant": (4007383, 62450, 740737, 50), "consph": (6010480, 83333, 214464, 20), "cop20k_A": (2624331, 121191, 104806, 73),
    "dense2": (4000000, 1999, 618114, 10), "jonheart": (37035, 3128, 684279, 60), "mac_econ_fwd500": (1273389, 206499, 759983, 12),
    "mc2depi": (2100225, 525824, 10765, 70), "pdb1HYS": (4344765, 36416, 426573, 30), "pwtk": (11634424, 217917, 281408, 25),
    "qcd5_4": (1916928, 49151, 17216, 63), "rail4284": (11279748, 4283, 560567, 10), "rma10": (2374001, 46834, 802446, 74),
    "scircuit": (958936, 170997, 321417, 30), "shipsec1": (7813404, 140873, 693767, 10), "webbase-1M": (3105536, 1000004, 223734, 77),
}
GTX580_MS = {"cant": 107.911, "consph": 62.1167, "cop20k_A": 146.603, "dense2": 26.6759, "jonheart": 2.66506, "mac_econ_fwd500": 23.5443,
             "mc2depi": 302.068, "pdb1HYS": 63.5818, "pwtk": 158.697, "qcd5_4": 74.6078, "rail4284": 58.7522, "rma10": 99.9466,
             "scircuit": 47.8156, "shipsec1": 42.4012, "webbase-1M": 592.115}   # the student's final numbers, ...tex:237-251


def write_vec(f, v, fmt):
    # chunked text writer: np.savetxt is too slow for 11 M values
    for i in range(0, len(v), 1 << 20):
        f.write(" ".join(fmt % t for t in v[i:i + (1 << 20)]))
        f.write(" ")
    f.write("\n")


def run(cmd, cwd, timeout=900):
    t0 = time.time()
    r = subprocess.run(cmd, cwd=cwd, capture_output=True, text=True, timeout=timeout)
    return r, time.time() - t0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    ap.add_argument("--iters-cap", type=int, default=0, help="cap N (0 = the published N)")
    args = ap.parse_args()
    import oracle
    from oracle import generators as g
    print("# The reference's own GPU programs on this B200 beside the drop-in drivers\n")
    print("Machine: %s" % subprocess.run(["nvidia-smi", "--query-gpu=name", "--format=csv,noheader"], capture_output=True, text=True).stdout.strip())
    # ---- PageRank -------------------------------------------------------------------------------------------
    with tempfile.TemporaryDirectory() as d:
        rows = []
        for name, cmd in (("reference pagerank.cu (sm_100a build)", [os.path.join(REF, "pagerank_ref")]),
                          ("bin/pagerank (this library)", [os.path.join(BIN, "pagerank")])):
            if not os.path.exists(cmd[0]):
                rows.append((name, "not built", ""))
                continue
            r, _ = run(cmd, d)
            m = re.search(r"gpu graph propagate took ([0-9.eE+-]+) ms", r.stdout)
            ok = "match" in r.stdout.lower() and "not" not in r.stdout.lower().split("match")[0][-20:]
            rows.append((name, m.group(1) if m else "?", "check passed" if ok else (r.stdout.strip().splitlines()[-1] if r.stdout.strip() else "rc %d" % r.returncode)))
        print("\n## PageRank, N = 2^21, avg_edges = 8, 20 iterations (hw/hw1/programming/pagerank.cu:149-151)\n")
        print("| program | gpu graph propagate, ms | its own 10-ULP check |\n|---|---|---|")
        for row in rows:
            print("| %s | %s | %s |" % row)
        print("\nGTX-era figure for the same run: 1188.11 ms (hw/hw1/programming/analysis/bandwidth_vs_avg_edges.csv).")
    # ---- final project -------------------------------------------------------------------------------------
    print("\n## Final project, the 15 benchmark shapes with seeded readMM.py-style inputs\n")
    print("| input | n | p | q | N | reference fp.cu ms | its rel-L2 vs double | bin/fp ms | its rel-L2 vs double | speed-up | GTX 580 ms (published) |")
    print("|---|---|---|---|---|---|---|---|---|---|---|")
    only = set(x for x in args.only.split(",") if x)
    for name, (n, p, q, N) in SHAPES.items():
        if only and name not in only:
            continue
        if args.iters_cap:
            N = min(N, args.iters_cap)
        a, s, k, x = g.fp_inputs(n, p, q, seed=sum(map(ord, name)), dtype=np.float32)
        a *= 0.999                                     # |a x| < 1: N products stay finite
        want = oracle.sparse_mv_scan(a.astype(np.float64), s, k, x.astype(np.float64), N)
        with tempfile.TemporaryDirectory() as d:
            with open(os.path.join(d, "a.txt"), "w") as f:
                f.write("%d %d %d %d\n" % (n, len(s), q, N))
                write_vec(f, a, "%.9g"); write_vec(f, s, "%d"); write_vec(f, k, "%d")
            with open(os.path.join(d, "x.txt"), "w") as f:
                write_vec(f, x, "%.9g")
            res = {}
            for key, exe in (("ref", os.path.join(REF, "fp_ref")), ("ours", os.path.join(BIN, "fp"))):
                if not os.path.exists(exe):
                    res[key] = ("not built", "")
                    continue
                try:
                    r, _ = run([exe, "a.txt", "x.txt"], d)
                except subprocess.TimeoutExpired:
                    res[key] = ("timeout", "")
                    continue
                m = re.search(r"iterations is: ([0-9.eE+-]+) milliseconds", r.stdout)
                err = ""
                bpath = os.path.join(d, "b.txt")
                if os.path.exists(bpath):
                    b = np.array(open(bpath).read().split(), dtype=np.float64)
                    os.remove(bpath)
                    if len(b) == n:
                        nb = np.nan_to_num(b, nan=1e30, posinf=1e30, neginf=-1e30)
                        err = "%.2e" % (np.linalg.norm(nb - want) / max(np.linalg.norm(want), 1e-300))
                res[key] = (m.group(1) if m else "rc %d" % r.returncode, err)
        try:
            sp = "%.1fx" % (float(res["ref"][0]) / float(res["ours"][0]))
        except Exception:
            sp = ""
        print("| %s | %d | %d | %d | %d | %s | %s | %s | %s | %s | %s |" % (name, n, len(s), q, N, res["ref"][0], res["ref"][1], res["ours"][0], res["ours"][1], sp,
                                                                           GTX580_MS.get(name, "")), flush=True)
    print("\nBoth programs time only their iteration loop (fp.cu:168-185; the drop-in keeps the x gather inside it, the reference "
          "hoists it to the host, fp.cu:124-125).  The reference kernel is warp-synchronous Fermi code without volatile or "
          "__syncwarp (fp.cu:28-59): where its rel-L2 column is far from 1e-6 it computed wrong sums on this GPU.")


if __name__ == "__main__":
    main()
