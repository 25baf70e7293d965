"""Where do the 1.26 GB of DRAM reads above the algorithmic bytes of the C5 product come from?  Runs the CSR kernel on
the 512^3 Laplacian three times: as it is; with every column index replaced by the row index (no x reuse across rows at
all: if the extra reads stay, they are not x); with the x vector only 4 MB long (columns mod 2^19: x always L2-resident).
Run plain for the times, or under `ncu --metrics dram__bytes_read.sum,... -k regex:csr_stream_lean` for the bytes."""
import ctypes as C
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("2012-04_stanford_cme213_b200")


def main():
    g = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    rows = g * g * g
    ctx = pkg.Context(0)
    nnz = C.c_int64()
    d_ptr = torch.empty(rows + 1, dtype=torch.int32, device="cuda")
    ctx.call("cme213_gen_laplace3d_f64", g, g, g, 0, rows, d_ptr, None, None, C.byref(nnz))
    d_col = torch.empty(nnz.value, dtype=torch.int32, device="cuda")
    d_val = torch.empty(nnz.value, dtype=torch.float64, device="cuda")
    ctx.call("cme213_gen_laplace3d_f64", g, g, g, 0, rows, d_ptr, d_col, d_val, C.byref(nnz))
    x = torch.rand(rows, dtype=torch.float64, device="cuda")
    y = torch.empty_like(x)
    counts = (d_ptr[1:] - d_ptr[:-1]).to(torch.int64)
    col_diag = torch.repeat_interleave(torch.arange(rows, dtype=torch.int32, device="cuda"), counts)
    col_small = torch.bitwise_and(d_col, (1 << 19) - 1)
    del counts
    torch.cuda.synchronize()

    def run(name, col):
        ctx.lib.cme213_ctx_clear_plans(ctx.h)
        for _ in range(3):
            ctx.call("cme213_spmv_csr_f64", 0, rows, d_ptr, col, d_val, x, y, 3, stream=torch.cuda.current_stream())
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            ctx.call("cme213_spmv_csr_f64", 0, rows, d_ptr, col, d_val, x, y, 3, stream=torch.cuda.current_stream())
        e1.record()
        torch.cuda.synchronize()
        print("%-28s %.4f ms per product" % (name, e0.elapsed_time(e1) / 5), flush=True)

    run("stencil columns", d_col)
    run("column = row", col_diag)
    run("columns mod 2^19 (4 MB of x)", col_small)


if __name__ == "__main__":
    main()
