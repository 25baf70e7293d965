"""fp32 5-point grid n x n: what the dispatcher picks, and every stream kernel id timed on the same box (cold L2: 8 copies of the
matrix in rotation for n <= 1024, one matrix larger than L2 otherwise).  python scripts/exp_c1_kernels.py [n ...]"""
import ctypes as C
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("2012-04_stanford_cme213_b200")


def main():
    ctx = pkg.Context(0)
    lib = ctx.lib
    for n in [int(a) for a in sys.argv[1:]] or [1024, 4096]:
        rows = n * n
        copies = 8 if n <= 1024 else 1
        mats = []
        nnz = C.c_int64()
        for _ in range(copies):
            p = torch.empty(rows + 1, dtype=torch.int32, device="cuda")
            ctx.call("cme213_gen_laplace2d_f32", n, n, p, None, None, C.byref(nnz))
            c = torch.empty(nnz.value, dtype=torch.int32, device="cuda")
            v = torch.empty(nnz.value, dtype=torch.float32, device="cuda")
            ctx.call("cme213_gen_laplace2d_f32", n, n, p, c, v, C.byref(nnz))
            mats.append((p, c, v, torch.rand(rows, device="cuda"), torch.empty(rows, device="cuda")))
        plan = C.c_void_p()
        rc = lib.cme213_csr_plan_create(ctx.h, 0, rows, rows, C.c_void_p(mats[0][0].data_ptr()), 4, None, C.byref(plan))
        print("n = %d: plan rc %d kernel %d: %s" % (n, rc, lib.cme213_csr_plan_kernel(plan), lib.cme213_csr_plan_describe(plan).decode()), flush=True)
        lib.cme213_csr_plan_destroy(plan)
        alg = nnz.value * 8 + (rows + 1) * 4 + rows * 8
        for kid, name in ((3, "stream (register-fed)"), (7, "stream-bulk"), (7 | 256, "stream-pipe"), (0, "auto")):
            lib.cme213_ctx_clear_plans(ctx.h)

            def one_round():
                for p, c, v, x, y in mats:
                    ctx.call("cme213_spmv_csr_f32", 0, rows, p, c, v, x, y, kid, stream=torch.cuda.current_stream())
            one_round()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                one_round()
            for _ in range(3):
                g.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 20 / copies
            print("  kernel %3d %-22s %.5f ms per product  %.0f GB/s" % (kid, name, ms, alg / ms / 1e6), flush=True)
            del g
        del mats
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
