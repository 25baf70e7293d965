"""CSR kernel sweep on one B200: scalar / vector(G) / stream / stream-padded / AUTO on matrices of mean row length
2, 8, 16, 64, 256 (regular) and the power-law matrix of config C4, with uniform-random and banded columns.  The table is
what the dispatcher's thresholds (csrc/spmv_csr.cu, cme213_csr_plan_*) were read from and what checks AUTO against the
best column.  Usage (GPU box): python scripts/sweep_csr.py [--rows-log2 22] > profiles/r02_csr_kernel_sweep.csv"""
import argparse
import ctypes as C
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nnz-log2", type=int, default=27, help="nonzeros per matrix (rows = nnz / mean length)")
    ap.add_argument("--reps", type=int, default=10)
    args = ap.parse_args()
    import torch
    pkg = importlib.import_module("2012-04_stanford_cme213_b200")
    ctx = pkg.Context(0)
    lib = ctx.lib
    has_plan = hasattr(lib, "cme213_csr_plan_create")
    kernels = [("scalar", 1), ("vec2", 2 | 2 << 8), ("vec4", 2 | 4 << 8), ("vec8", 2 | 8 << 8), ("vec16", 2 | 16 << 8), ("vec32", 2 | 32 << 8),
               ("stream", 3), ("padded", 4), ("auto", 0)]
    print("matrix,rows,nnz,window," + ",".join(k for k, _ in kernels) + ",best,auto_vs_best")
    cases = [("len%d" % L, L, False) for L in (2, 8, 16, 64, 256)] + [("powerlaw", 0, True)]
    for name, L, power in cases:
        for window in (0, 4096):
            rows = (1 << 24) if power else (1 << args.nnz_log2) // L
            d_ptr = torch.empty(rows + 1, dtype=torch.int32, device="cuda")
            nnz = C.c_int64()
            if power:
                ctx.call("cme213_gen_powerlaw_lengths", rows, 4, 265, 100, 256, 6, d_ptr, C.byref(nnz))
            else:
                ctx.call("cme213_gen_powerlaw_lengths", rows, L, 0, 1, L, 6, d_ptr, C.byref(nnz))
            d_col = torch.empty(nnz.value, dtype=torch.int32, device="cuda")
            d_val = torch.empty(nnz.value, dtype=torch.float32, device="cuda")
            ctx.call("cme213_gen_random_csr_fill_f32", rows, rows, d_ptr, window, 7, d_col, d_val)
            x = torch.empty(rows, dtype=torch.float32, device="cuda")
            ctx.call("cme213_fill_uniform_f32", x, rows, 8)
            y = torch.empty_like(x)
            alg = nnz.value * 8 + (rows + 1) * 4 + rows * 8
            times = []
            ctx.lib.cme213_ctx_clear_plans(ctx.h)     # AUTO's plans are keyed by address; the allocator re-uses addresses between matrices
            for kname, kid in kernels:
                try:
                    for _ in range(3):
                        ctx.call("cme213_spmv_csr_f32", 0, rows, d_ptr, d_col, d_val, x, y, kid)
                    torch.cuda.synchronize()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    for _ in range(args.reps):
                        ctx.call("cme213_spmv_csr_f32", 0, rows, d_ptr, d_col, d_val, x, y, kid)
                    e1.record()
                    torch.cuda.synchronize()
                    times.append(e0.elapsed_time(e1) / args.reps)
                except Exception as ex:  # noqa: BLE001
                    print("# %s %s failed: %r" % (name, kname, ex), file=sys.stderr)
                    times.append(float("nan"))
            best = min(t for t in times[:-1] if t == t)
            print("%s,%d,%d,%d,%s,%.4f,%.3f" % (name, rows, nnz.value, window, ",".join("%.4f" % t for t in times), best, times[-1] / best), flush=True)
            print("# %s window %d: best %.3f ms = %.0f GB/s algorithmic" % (name, window, best, alg / best / 1e6), file=sys.stderr)
            del d_ptr, d_col, d_val, x, y
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
