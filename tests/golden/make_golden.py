"""Generates tests/golden/*.npz by running the REFERENCE's own CPU code (oracle/_ref, compiled from
/root/reference by oracle/Makefile) on small seeded inputs, and copies the reference's known-answer
fixture.  Run in the authoring container only (it needs /root/reference):

    python tests/golden/make_golden.py

The outputs are committed; the GPU box never sees /root/reference.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from oracle import generators as g  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def main():
    oracle.build()
    assert oracle.ref_lib("pagerank") is not None and oracle.ref_lib("spmvscan") is not None, "build oracle/_ref first"

    # 1. the reference's known-answer fixture (hw/hw_final/programming/aux/CheckOutput/small_*.txt)
    d = os.path.join(REF, "hw/hw_final/programming/aux/CheckOutput")
    tok = open(os.path.join(d, "small_a.txt")).read().split()
    n, p, q, N = (int(t) for t in tok[:4])
    a = np.array([float(t) for t in tok[4:4 + n]])
    s = np.array([int(t) for t in tok[4 + n:4 + n + p]], dtype=np.int32)
    k = np.array([int(t) for t in tok[4 + n + p:4 + n + p + n]], dtype=np.int32)
    x = np.array([float(t) for t in open(os.path.join(d, "small_x.txt")).read().split()])
    b = np.array([float(t) for t in open(os.path.join(d, "small_b.txt")).read().split()])
    np.savez(os.path.join(HERE, "fp_small.npz"), n=n, p=p, q=q, N=N, a=a, s=s, k=k, x=x, b=b)

    # 2. sparseMV_scan (instructor reference, double) on seeded inputs, 1 and 3 iterations
    a, s, k, x = g.fp_inputs(6000, 400, 257, seed=11, dtype=np.float64)
    np.savez(os.path.join(HERE, "fp_ref_n6000.npz"), a=a, s=s, k=k, x=x,
             b1=oracle.ref_sparse_mv_scan(a, s, k, x, 1), b3=oracle.ref_sparse_mv_scan(a, s, k, x, 3))
    # one long segment + many length-1 segments (edge cases of the loader's invariants)
    a, _, k, x = g.fp_inputs(3000, 2, 31, seed=12, dtype=np.float64)
    s = np.concatenate([[0], np.arange(2000, 3001)]).astype(np.int32)
    np.savez(os.path.join(HERE, "fp_ref_skewed.npz"), a=a, s=s, k=k, x=x, b2=oracle.ref_sparse_mv_scan(a, s, k, x, 2))

    # 3. host_graph_iterate (pagerank.cu:59-67) on the homework's graph recipe with a seeded hash
    ind, edges, inv, nodes = g.pagerank_graph(5000, 8, seed=21)
    np.savez(os.path.join(HERE, "pagerank_ref_n5000.npz"), indices=ind, edges=edges, inv=inv, nodes=nodes,
             out20=oracle.ref_pagerank_iterate(ind, edges, nodes, inv, 20), out2=oracle.ref_pagerank_iterate(ind, edges, nodes, inv, 2))

    # 4. Bell & Garland SC09 Fig. 2 / Fig. 4 example (4x4 matrix) -- typed in from the paper, '*' padding = 0
    np.savez(os.path.join(HERE, "bell_fig2.npz"),
             csr_ptr=np.array([0, 2, 4, 7, 9], np.int32), csr_idx=np.array([0, 1, 1, 2, 0, 2, 3, 1, 3], np.int32),
             csr_val=np.array([1, 7, 2, 8, 5, 3, 9, 6, 4], np.float32),
             ell_val_colmajor=np.array([1, 2, 5, 6, 7, 8, 3, 4, 0, 0, 9, 0], np.float32),
             ell_idx_rows=np.array([[0, 1, -1], [1, 2, -1], [0, 2, 3], [1, 3, -1]], np.int32),
             coo_row=np.array([0, 0, 1, 1, 2, 2, 2, 3, 3], np.int32))
    # 5. the reference's radix sort passes (hw/hw4/programming/radixsort.cpp) on 12 000 seeded keys: 31-bit values
    #    like its rand() input (:173-175) with a low-entropy stretch so that stability matters
    assert oracle.ref_lib("radixsort") is not None
    rng = np.random.default_rng(2012)
    keys = rng.integers(0, 2**31, 12000, dtype=np.int64).astype(np.uint32)
    keys[3000:6000] = (keys[3000:6000] & 0x00030003) | 0x5A000000
    p0 = oracle.ref_radix_pass(keys, 0, 8)
    p1 = oracle.ref_radix_pass(p0, 8, 4, block_size=1000)
    p2 = oracle.ref_radix_pass(keys, 13, 5, block_size=0)
    s8, _ = oracle.ref_radix_sort(keys, 8)
    s4, _ = oracle.ref_radix_sort(keys, 4)
    np.savez_compressed(os.path.join(HERE, "radixsort_ref_n12000.npz"), keys=keys, pass_0_8=p0, pass_8_4_of_pass_0_8=p1,
                        serial_pass_13_5=p2, sorted_bits8=s8, sorted_bits4=s4)
    print("golden vectors written to", HERE)


if __name__ == "__main__":
    main()
