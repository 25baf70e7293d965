"""Reproduces the homework's bandwidth-vs-avg_edges sweep (hw/hw1/programming/analysis/pagerank.cu,
analysis/bandwidth_vs_avg_edges.csv: 6.99 -> 4.01 GB/s for avg_edges 2 -> 20 on the student's GPU) on B200 through
cme213_pagerank_propagate.  N = 2^21 nodes, 20 iterations per point, bytes counted like the homework does
(edges once, indices, and three N-vectors per iteration).  Writes CSV to stdout."""
import ctypes as C
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("2012-04_stanford_cme213_b200")


def main():
    torch.cuda.set_device(0)
    ctx = pkg.Context(0)
    n, iters = 1 << 21, 20
    print("avg_edges,nnz,ms_per_iteration,GB_per_s")
    for avg in range(2, 21):
        ind = torch.empty(n + 1, dtype=torch.int32, device="cuda")
        inv = torch.empty(n, dtype=torch.float32, device="cuda")
        a = torch.empty(n, dtype=torch.float32, device="cuda")
        nnz = C.c_int64()
        ctx.call("cme213_gen_pagerank_graph", n, avg, 21, ind, None, inv, a, C.byref(nnz))
        edges = torch.empty(nnz.value, dtype=torch.int32, device="cuda")
        ctx.call("cme213_gen_pagerank_graph", n, avg, 21, ind, edges, inv, a, C.byref(nnz))
        b = torch.empty_like(a)
        for _ in range(4):
            ctx.call("cme213_pagerank_propagate", ind, edges, a, b, inv, n, 0)
            ctx.call("cme213_pagerank_propagate", ind, edges, b, a, inv, n, 0)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters // 2):
            ctx.call("cme213_pagerank_propagate", ind, edges, a, b, inv, n, 0)
            ctx.call("cme213_pagerank_propagate", ind, edges, b, a, inv, n, 0)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        bytes_it = nnz.value * 4 + (n + 1) * 4 + n * 12
        print("%d,%d,%.4f,%.1f" % (avg, nnz.value, ms, bytes_it / ms / 1e6))


if __name__ == "__main__":
    main()
