#!/usr/bin/env python
"""bench.py -- measures the hot path on B200 (see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c5|c1|scan|segscan|radixsort|dia|stencil|fp|transpose32|
                     transpose64|ell|hyb|csr_powerlaw|pagerank] [--impl ours|reference] [--extras 0|1]

Default workload = BASELINE.json config C5: row-sharded CSR SpMV iteration on the 3-D 7-point
Laplacian of a 512^3 grid in fp64 (134,217,728 rows, 937,951,232 nonzeros).  A *step* is one pass
y = A x over the whole matrix, including the x exchange when N > 1 (strong scaling: the matrix is
fixed, rows are split across ranks).  `value` is aggregate algorithmic GB/s: the compulsory bytes of
one SpMV (matrix arrays once, x once, y once) times K, divided by the max-over-ranks device time.

One JSON line on stdout (rank 0).  Everything else goes to stderr.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "2012-04_stanford_cme213_b200"


# roofline.traffic is READ from the committed ncu summary of the dominant kernel (made by scripts/ncu_summary.py from an
# `ncu --set full` capture of this command), never typed in: {workload: (file under profiles/, kernel-name fragment)}.
NCU_SUMMARY = {"c5": ("ncu_c5_csr_kernel.txt", "csr_stream_lean_kernel<double")}


def ncu_traffic(workload):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch, from profiles/<file>; None when the file is absent,
    names another kernel or cannot be parsed (then the bench line says traffic: null rather than a stale figure)."""
    if workload not in NCU_SUMMARY:
        return None, None
    fname, frag = NCU_SUMMARY[workload]
    path = os.path.join(ROOT, "profiles", fname)
    scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12}
    try:
        total, seen_kernel = 0.0, False
        with open(path) as f:
            for line in f:
                t = line.split()
                if line.startswith("kernel:"):
                    if seen_kernel:
                        break               # first kernel block only
                    if frag not in line:
                        return None, None
                    seen_kernel = True
                elif seen_kernel and t and t[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                    total += float(t[1]) * scale[t[2]]
        return (total, "profiles/" + fname) if total > 0 else (None, None)
    except Exception:
        return None, None


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            f = [t.strip() for t in r.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------
# workloads: each returns dict(step=callable, alg_bytes, flops, units, kernel, launches_per_step, meta)
# ---------------------------------------------------------------------------------------------------
def i32(n, torch):
    return torch.empty(int(n), dtype=torch.int32, device="cuda")


def make_vectors(sh, ctx, torch, dist, rank, world, rows, halo, exchange):
    """The two full-length ping-pong vectors.  `fused` needs vectors the neighbours can map (one cudaMalloc each,
    exported through CUDA IPC) and shard / halo sizes that are whole 256-row blocks; if any rank cannot set that
    up, every rank falls back to `p2p` (NCCL send/recv)."""
    if world > 1 and exchange == "fused":
        ok, xa = 1, None
        try:
            per = rows // world
            if rows % world or per % 256 or halo % 256 or per < 2 * halo:
                raise ValueError("shape not suited to the fused exchange")
            xa, pa = sh.shared_vector(ctx.lib, rows)
            xb, pb = sh.shared_vector(ctx.lib, rows)
            xb.zero_()
        except Exception as e:  # noqa: BLE001
            log("[bench] fused exchange unavailable on rank %d (%r)" % (rank, e))
            ok = 0
        flag = torch.tensor([ok], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 1:
            peers = None
            try:
                peers = sh.PeerVectors(ctx.lib, dist, rank, world, [pa, pb])
            except Exception as e:  # noqa: BLE001 -- IPC refused by the driver / container
                log("[bench] CUDA IPC unavailable on rank %d (%r)" % (rank, e))
                ok = 0
            flag = torch.tensor([ok], device="cuda")
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)    # every rank reaches the same conclusion
            if int(flag.item()) == 1:
                return xa, xb, peers, "fused"
        exchange = "p2p"
    xa = torch.empty(rows, dtype=torch.float64, device="cuda")
    xb = torch.zeros(rows, dtype=torch.float64, device="cuda")
    return xa, xb, None, exchange


def make_c5_shard(ctx, torch, args, rank, world, dist):
    """N > 1 through the C ABI's own sharded iterate (csrc/sharded.cu): one kernel launch per pass, halo rows stored into
    the neighbours' vectors by that kernel, passes ordered by flags in peer memory -- no NCCL on the data path.  Python
    only carries the export handles between the ranks and replays one CUDA graph per pass direction."""
    import ctypes as C
    sh = importlib.import_module(PKG + ".sharded")
    g = args.grid
    rows = g * g * g
    bounds = sh.partition_rows(ctx.lib, rows, world, None, 256)       # the stencil's nonzeros are spread evenly: equal rows
    lo, hi = bounds[rank], bounds[rank + 1]
    nloc = hi - lo
    nnz = C.c_int64()
    d_ptr = i32(nloc + 1, torch)
    ctx.call("cme213_gen_laplace3d_f64", g, g, g, lo, hi, d_ptr, None, None, C.byref(nnz))
    d_col = i32(nnz.value, torch)
    d_val = torch.empty(nnz.value, dtype=torch.float64, device="cuda")
    ctx.call("cme213_gen_laplace3d_f64", g, g, g, lo, hi, d_ptr, d_col, d_val, C.byref(nnz))
    shard = sh.Shard(ctx, dist, rank, world, rows, bounds, d_ptr, d_col, d_val, mode={"flags": 0, "gather": 2, "split": 3}[args.exchange])
    xa, xb = shard.x
    ctx.call("cme213_fill_uniform_f64", xa, rows, 9)
    xa.mul_(0.05)
    xb.zero_()
    torch.cuda.synchronize()
    dist.barrier()
    state = {"src": 0}
    graphs = None
    if args.graph and shard.mode != 3:
        shard.step(0); shard.step(1)             # eager pair first: kernels loaded, nothing left to set up inside the capture
        shard.finish()
        dist.barrier()
        graphs = []
        for src in (0, 1):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                shard.step(src)
            graphs.append(gr)
        # the capture itself launches nothing, but keep every rank's pass counters in step: nothing to do
        torch.cuda.synchronize()
        dist.barrier()

    def step():
        if graphs is not None:
            graphs[state["src"]].replay()
        else:
            shard.step(state["src"])
        state["src"] = 1 - state["src"]

    def kernel_step():   # the SpMV kernel alone on this rank's rows (roofline leg)
        ctx.call("cme213_spmv_csr_f64", 0, nloc, d_ptr, d_col, d_val, xa, xb.data_ptr() + lo * 8, args.csr_kernel,
                 stream=torch.cuda.current_stream())

    modes = {1: "flags (halo rows stored into the neighbours' vectors by the SpMV kernel, ordered by flags in peer memory)",
             2: "gather (peer copies of the whole slice to every rank)", 3: "gather_split (local-column block overlapped with the peer copies)"}
    nnz_total = 7 * rows - 6 * g * g
    alg = nnz_total * 12 + (rows + 1) * 4 + rows * 8 * 2
    local_alg = nnz.value * 12 + (nloc + 1) * 4 + nloc * 8 * 2
    return dict(step=step, alg_bytes=alg, local_alg_bytes=local_alg, flops=2 * nnz_total, kernel="csr_stream_lean_kernel<double>",
                launches_per_step=1 if shard.mode == 1 else 3, kernel_step=kernel_step,
                meta={"workload": C5_WORKLOAD % g, "sharding": "rows split over %d GPU(s)" % world,
                      "rows": rows, "nnz": nnz_total, "exchange": modes[shard.mode], "comm_bytes_per_rank_per_step": shard.comm_bytes_per_step,
                      "launch": "one CUDA graph per pass direction (1 kernel node each); C ABI cme213_shard_step" if graphs is not None else "eager cme213_shard_step",
                      "cta_order": "boundary row blocks first, then row order",
                      "l2": "working set %.1f GB per pass >> 126 MB L2" % (alg / 1e9)},
                keep=(d_ptr, d_col, d_val, graphs), shard_obj=shard, state=state, shard=(lo, hi, d_ptr, d_col, d_val, nnz.value), ctx=ctx)


def make_c5(ctx, torch, args, rank, world, dist):
    import ctypes as C
    pkg = importlib.import_module(PKG)
    sh = pkg.sharded if hasattr(pkg, "sharded") else importlib.import_module(PKG + ".sharded")
    if world > 1 and args.exchange in ("flags", "gather", "split"):
        try:
            return make_c5_shard(ctx, torch, args, rank, world, dist)
        except RuntimeError as e:   # raised on every rank together (Shard gathers the outcome): e.g. CUDA IPC refused
            log("[bench] rank %d: C-ABI sharded iterate unavailable (%s); falling back to the NCCL-ordered fused exchange" % (rank, e))
            args.exchange = "fused"
    g = args.grid
    rows = g * g * g
    lo, hi = sh.row_partition(rows, world)[rank]
    nloc = hi - lo
    nnz = C.c_int64()
    d_ptr = i32(nloc + 1, torch)
    ctx.call("cme213_gen_laplace3d_f64", g, g, g, lo, hi, d_ptr, None, None, C.byref(nnz))
    d_col = i32(nnz.value, torch)
    d_val = torch.empty(nnz.value, dtype=torch.float64, device="cuda")
    ctx.call("cme213_gen_laplace3d_f64", g, g, g, lo, hi, d_ptr, d_col, d_val, C.byref(nnz))
    xa, xb, peers, exchange = make_vectors(sh, ctx, torch, dist, rank, world, rows, g * g, args.exchange)
    ctx.call("cme213_fill_uniform_f64", xa, rows, 9)
    xa.mul_(0.05)  # keeps 7-point growth (|lambda| <= 12) far from overflow over the run
    streams = (torch.cuda.current_stream(), torch.cuda.Stream()) if world > 1 else None
    order_fn = (lambda a, b: sh.grid3d_block_order(g, g, g, a, b, strip_lines=args.order_strip)) if args.order_strip > 0 else None
    it = sh.ShardedIterate(rows, world, rank, sh.cuda_spmv(ctx, lo, d_ptr, d_col, d_val, args.csr_kernel, order_fn), xa, xb,
                           halo=g * g, mode=exchange, dist=dist, streams=streams, peers=peers)
    state = {"src": 0}
    replay = it.capture_pair() if world > 1 else None

    def step():
        if replay is not None:
            if state["src"] == 0:
                replay()          # one graph launch = two passes (A->B, B->A); the odd call is already done
            state["src"] = 1 - state["src"]
            return
        it.step(state["src"], 1 - state["src"])
        state["src"] = 1 - state["src"]

    nnz_total = 7 * rows - 6 * g * g
    alg = nnz_total * 12 + (rows + 1) * 4 + rows * 8 * 2
    local_alg = nnz.value * 12 + (nloc + 1) * 4 + nloc * 8 * 2
    kernel_step = lambda: it.spmv(it.lo, it.hi, it.x[0], it.x[1])
    return dict(step=step, alg_bytes=alg, local_alg_bytes=local_alg, flops=2 * nnz_total, kernel="csr_stream_lean_kernel<double>",
                launches_per_step=1 if it.mode not in ("halo", "p2p", "fused") else 2, kernel_step=kernel_step, state=state,
                meta={"workload": C5_WORKLOAD % g, "sharding": "rows split over %d GPU(s)" % world,
                      "rows": rows, "nnz": nnz_total, "exchange": it.mode, "comm_bytes_per_rank_per_step": it.comm_bytes_per_step,
                      "launch": ("CUDA graph of two passes (kernels + NCCL %s)" % ({"p2p": "send/recv", "fused": "1-element all-reduce; halo pushed by the kernel"}.get(it.mode, "all-gather"))) if replay is not None else "eager",
                      "cta_order": ("z fastest inside strips of %d grid lines" % args.order_strip) if args.order_strip > 0 else "row order",
                      "l2": "working set %.1f GB per pass >> 126 MB L2" % (alg / 1e9)},
                keep=(d_ptr, d_col, d_val, xa, xb, it), iterate=it, shard=(lo, hi, d_ptr, d_col, d_val, nnz.value), ctx=ctx)


def make_c1(ctx, torch, args):
    import ctypes as C
    n = args.c1_n
    rows = n * n
    copies = 8 if n >= 1024 else min(64, 8 * (1024 // n) ** 2)  # 8 x 54 MB > 126 MB L2: rotate so every step streams from HBM
    mats = []
    nnz = C.c_int64()
    for _ in range(copies):
        d_ptr = i32(rows + 1, torch)
        ctx.call("cme213_gen_laplace2d_f32", n, n, d_ptr, None, None, C.byref(nnz))
        d_col, d_val = i32(nnz.value, torch), torch.empty(nnz.value, dtype=torch.float32, device="cuda")
        ctx.call("cme213_gen_laplace2d_f32", n, n, d_ptr, d_col, d_val, C.byref(nnz))
        x = torch.empty(rows, dtype=torch.float32, device="cuda")
        ctx.call("cme213_fill_uniform_f32", x, rows, 2)
        mats.append((d_ptr, d_col, d_val, x, torch.empty_like(x)))
    def one_round():
        for p, c, v, x, y in mats:
            ctx.call("cme213_spmv_csr_f32", 0, rows, p, c, v, x, y, args.csr_kernel, stream=torch.cuda.current_stream())

    # An 8 us kernel cannot be fed from Python one launch at a time (the host needs ~25 us per call): one step is a
    # CUDA graph of the 8 rotating copies, i.e. 8 back-to-back passes, and the per-pass time is the step over 8.
    one_round()                      # warm-up outside the capture (kernel choice is cached per matrix)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        one_round()
    torch.cuda.synchronize()

    alg = nnz.value * 8 + (rows + 1) * 4 + rows * 8

    # What this size allows at all: a plain device copy moving the same bytes per launch (half read, half written), launched
    # the same way (graph of 8 back-to-back launches over rotating buffers, cold L2).  A probe for the notes, not a product path.
    half = (alg // 2 + 15) // 16 * 16
    bufs = [(torch.empty(half, dtype=torch.uint8, device="cuda"), torch.empty(half, dtype=torch.uint8, device="cuda")) for _ in range(copies)]
    for a, b in bufs:
        b.copy_(a)
    torch.cuda.synchronize()
    cgraph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(cgraph):
        for a, b in bufs:
            b.copy_(a)
    copy_ms = time_steps(torch, cgraph.replay, 20, 3) / 20 / copies
    del cgraph, bufs
    return dict(step=graph.replay, alg_bytes=alg * copies, flops=2 * nnz.value * copies, kernel="csr_stream_pipe_kernel<float>" if args.csr_kernel in (0, 8, 263) else "csr_stream_bulk_kernel<float>" if args.csr_kernel == 7 else "csr_stream_lean_kernel<float>",
                launches_per_step=copies,
                meta={"workload": "C1: CSR SpMV fp32, 2-D 5-point Poisson %dx%d" % (n, n), "rows": rows, "nnz": nnz.value,
                      "passes_per_step": copies, "device_copy_of_the_same_bytes_ms_per_launch": copy_ms,
                      "l2": "one step = CUDA graph of %d passes over %d independent copies (%.0f MB total): every pass is cold in L2" % (copies, copies, copies * alg / 1e6)},
                keep=(mats, graph))


def make_scan(ctx, torch, args):
    n = 1 << 28
    x = i32(n, torch)
    ctx.call("cme213_fill_hash_u32", x, n, 3, 0)
    out = torch.empty_like(x)

    def step():
        ctx.call("cme213_exclusive_scan_u32", x, out, n, stream=torch.cuda.current_stream())

    return dict(step=step, alg_bytes=8 * n, flops=n, kernel="lookback_scan_kernel<u32>", launches_per_step=1,
                meta={"workload": "C2a: exclusive scan of 2^28 uint32 (full-range values, wraps)", "n": n,
                      "l2": "2.1 GB per pass >> L2"}, keep=(x, out))


def make_radixsort(ctx, torch, args):
    # 8f.4: LSD radix sort of 2^28 uint32 keys, 8-bit digits (the reference's default numBits, radixsort.cpp:166;
    # its default size is 40 M keys).  The four passes run out of place from a pristine buffer so that every
    # timed step sorts the same unsorted keys.
    n = 1 << 28
    x = i32(n, torch)
    ctx.call("cme213_fill_hash_u32", x, n, 5, 0)
    a, b = torch.empty_like(x), torch.empty_like(x)

    def step():
        st = torch.cuda.current_stream()
        ctx.call("cme213_radix_sort_pass_u32", x, a, n, 0, 8, stream=st)
        ctx.call("cme213_radix_sort_pass_u32", a, b, n, 8, 8, stream=st)
        ctx.call("cme213_radix_sort_pass_u32", b, a, n, 16, 8, stream=st)
        ctx.call("cme213_radix_sort_pass_u32", a, b, n, 24, 8, stream=st)

    return dict(step=step, alg_bytes=4 * 12 * n, flops=0, kernel="radix_scatter_kernel", launches_per_step=12,
                meta={"workload": "8f.4: LSD radix sort of 2^28 uint32 keys, 4 passes of 8 bits (12 B per key per pass: "
                                  "histogram read, scatter read + write)", "n": n,
                      "l2": "1.07 GB of keys >> L2"}, keep=(x, a, b))


def make_dia(ctx, torch, args):
    # 8f.4: config C5's operator (3-D 7-point Laplacian, fp64) in DIA, converted on the device from the CSR arrays
    import ctypes as C
    pkg = importlib.import_module(PKG)
    g = args.grid
    rows = g * g * g
    nnz = C.c_int64()
    d_ptr = i32(rows + 1, torch)
    ctx.call("cme213_gen_laplace3d_f64", g, g, g, 0, rows, d_ptr, None, None, C.byref(nnz))
    d_col = i32(nnz.value, torch)
    d_val = torch.empty(nnz.value, dtype=torch.float64, device="cuda")
    ctx.call("cme213_gen_laplace3d_f64", g, g, g, 0, rows, d_ptr, d_col, d_val, C.byref(nnz))
    d_off = i32(16, torch)
    nd = C.c_int()
    ctx.call("cme213_csr_diagonals", rows, rows, d_ptr, d_col, 16, d_off, C.byref(nd))
    pitch = int(pkg.load().cme213_ell_pitch(rows))
    d_dat = torch.empty(nd.value * pitch, dtype=torch.float64, device="cuda")
    miss = C.c_int64()
    ctx.call("cme213_csr_to_dia_f64", rows, d_ptr, d_col, d_val, nd.value, d_off, pitch, d_dat, C.byref(miss))
    assert miss.value == 0 and nd.value == 7
    del d_ptr, d_col, d_val
    torch.cuda.empty_cache()
    xa = torch.empty(rows, dtype=torch.float64, device="cuda")
    xb = torch.empty_like(xa)
    ctx.call("cme213_fill_uniform_f64", xa, rows, 9)
    xa.mul_(0.05)
    state = {"src": 0}

    def step():
        a, b = (xa, xb) if state["src"] == 0 else (xb, xa)
        ctx.call("cme213_spmv_dia_f64", rows, rows, pitch, 7, d_off, d_dat, a, b, stream=torch.cuda.current_stream())
        state["src"] = 1 - state["src"]

    csr_bytes = nnz.value * 12 + (rows + 1) * 4 + rows * 16
    return dict(step=step, alg_bytes=7 * rows * 8 + rows * 16, flops=2 * nnz.value, kernel="dia_v4_kernel<double>",
                launches_per_step=1,
                meta={"workload": "8f.4: C5's operator in DIA (7 diagonals, fp64, %d^3 grid): 72 B per row against CSR's 104" % g,
                      "rows": rows, "csr_bytes_for_the_same_product": csr_bytes, "l2": "9.7 GB per pass >> L2"},
                keep=(d_off, d_dat, xa, xb))


def make_stencil(ctx, torch, args):
    # 8f.4: config C5's operator with no matrix arrays at all (only x and y move): the upper bound for C5
    g = args.grid
    rows = g * g * g
    xa = torch.empty(rows, dtype=torch.float64, device="cuda")
    xb = torch.empty_like(xa)
    ctx.call("cme213_fill_uniform_f64", xa, rows, 9)
    xa.mul_(0.05)
    state = {"src": 0}

    def step():
        a, b = (xa, xb) if state["src"] == 0 else (xb, xa)
        ctx.call("cme213_stencil_laplace_f64", g, g, g, 6.0, 0, rows, a, b, stream=torch.cuda.current_stream())
        state["src"] = 1 - state["src"]

    nnz = 7 * rows - 6 * g * g
    return dict(step=step, alg_bytes=rows * 16, flops=2 * nnz, kernel="stencil_vec_kernel<double>", launches_per_step=1,
                meta={"workload": "8f.4: C5's operator matrix-free (7-point stencil, fp64, %d^3 grid): 16 B per row" % g,
                      "rows": rows, "csr_bytes_for_the_same_product": nnz * 12 + (rows + 1) * 4 + rows * 16,
                      "l2": "2.1 GB per pass >> L2"}, keep=(xa, xb))


def make_segscan(ctx, torch, args):
    n = 1 << 28
    v = torch.empty(n, dtype=torch.float32, device="cuda")
    ctx.call("cme213_fill_uniform_f32", v, n, 4)
    bits = i32((n + 31) // 32, torch)
    # segment offsets as readMM.py:44-49 makes them (sorted random cut points, n/64 of them), then the head-flag bitmask
    import ctypes as C
    cuts = n // 64
    d_s = i32(cuts + 2, torch)
    p = C.c_int64()
    ctx.call("cme213_gen_segment_offsets", n, cuts, 5, d_s, C.byref(p))
    ctx.call("cme213_segment_flags_from_offsets", d_s, int(p.value), n, bits)
    del d_s
    out = torch.empty_like(v)

    def step():
        ctx.call("cme213_segscan_inclusive_f32", v, bits, out, n, stream=torch.cuda.current_stream())

    return dict(step=step, alg_bytes=8 * n + n // 8, flops=n, kernel="lookback_scan_kernel<f32,seg>", launches_per_step=1,
                meta={"workload": "C2b: inclusive segmented fp32 scan, n=2^28, segments from n/64 sorted random cut points (readMM.py recipe, mean length 64)", "n": n,
                      "l2": "2.2 GB per pass >> L2"}, keep=(v, bits, out))


def make_fp(ctx, torch, args):
    n, q = 1 << 28, 1 << 20
    a = torch.empty(n, dtype=torch.float32, device="cuda")
    ctx.call("cme213_fill_uniform_f32", a, n, 4)
    a0 = a.clone()
    k = i32(n, torch)
    ctx.call("cme213_fill_hash_u32", k, n, 6, q)
    x = torch.empty(q, dtype=torch.float32, device="cuda")
    ctx.call("cme213_fill_uniform_f32", x, q, 7)
    bits = i32((n + 31) // 32, torch)
    ctx.call("cme213_fill_random_flags", bits, n, 5, (1 << 32) // 64)
    state = {"i": 0}

    def step():
        if state["i"] % 8 == 0:
            a.copy_(a0)  # values shrink by |x|<1 every pass; restart before they underflow (untimed share is small, noted)
        state["i"] += 1
        ctx.call("cme213_mul_gather_segscan_f32", a, k, x, bits, a, n, stream=torch.cuda.current_stream())

    return dict(step=step, alg_bytes=12 * n + 4 * q + n // 8, flops=2 * n, kernel="lookback_scan_kernel<mul_gather>", launches_per_step=1,
                meta={"workload": "C2c: final-project step b = segscan(a*x[k]), n=2^28, q=2^20, in place", "n": n,
                      "l2": "3.2 GB per pass >> L2", "note": "every 8th step restores a (one extra 2 GB copy inside the timed region)"},
                keep=(a, a0, k, x, bits))


def make_fp_iterate(ctx, torch, args):
    """C2c as the final project runs it: N = 10 iterations a <- b on the same x.  x[k] is gathered once (inside the timed
    step) and the ten iterations stream a and xx through the bulk pipeline; one bench step = the gather + 10 iterations."""
    n, q, N = 1 << 28, 1 << 20, 10
    a = torch.empty(n, dtype=torch.float32, device="cuda")
    ctx.call("cme213_fill_uniform_f32", a, n, 4)
    a0 = a.clone()
    k = i32(n, torch)
    ctx.call("cme213_fill_hash_u32", k, n, 6, q)
    x = torch.empty(q, dtype=torch.float32, device="cuda")
    ctx.call("cme213_fill_uniform_f32", x, q, 7)
    xx = torch.empty_like(a)
    bits = i32((n + 31) // 32, torch)
    ctx.call("cme213_fill_random_flags", bits, n, 5, (1 << 32) // 64)

    def step():
        st = torch.cuda.current_stream()
        ctx.call("cme213_gather_f32", x, k, xx, n, stream=st)
        ctx.call("cme213_mul_segscan_f32", a0, xx, bits, a, n, stream=st)       # first iteration reads the pristine a
        for _ in range(N - 1):
            ctx.call("cme213_mul_segscan_f32", a, xx, bits, a, n, stream=st)

    per_iter = 12 * n + n // 8
    return dict(step=step, alg_bytes=N * per_iter + 8 * n + 4 * q, flops=2 * n * N, kernel="scan_bulk_blocked_kernel<mul>", launches_per_step=N + 1,
                meta={"workload": "C2c iterated: final-project loop, N=10 iterations on n=2^28, q=2^20: x[k] gathered once, then a*xx + segscan streamed", "n": n,
                      "iterations_per_step": N, "l2": "3.2 GB per iteration >> L2",
                      "note": "ms is for the gather + 10 iterations; the fused single step (gather every iteration) is the C2c line above"},
                keep=(a, a0, k, x, xx, bits))


def make_transpose(ctx, torch, args, bits):
    n = 16384
    dt = torch.float32 if bits == 32 else torch.float64
    a = torch.empty((n, n), dtype=dt, device="cuda")
    ctx.call("cme213_fill_uniform_f32" if bits == 32 else "cme213_fill_uniform_f64", a, n * n, 8)
    b = torch.empty_like(a)
    fn = "cme213_transpose_b32" if bits == 32 else "cme213_transpose_b64"

    def step():
        ctx.call(fn, a, b, n, n, n, n, stream=torch.cuda.current_stream())

    return dict(step=step, alg_bytes=2 * n * n * (bits // 8), flops=0, kernel="transpose_vec_kernel<b%d>" % bits, launches_per_step=1,
                meta={"workload": "C3: out-of-place transpose 16384x16384 fp%d" % bits, "l2": "%.1f GB per pass >> L2" % (2 * n * n * bits / 8e9)},
                keep=(a, b))


def make_powerlaw(ctx, torch, args, fmt):
    import ctypes as C
    pkg = importlib.import_module(PKG)
    rows = args.powerlaw_rows
    d_ptr = i32(rows + 1, torch)
    nnz = C.c_int64()
    if args.regular:
        d_ptr.copy_((torch.arange(rows + 1, device="cuda", dtype=torch.int64) * 16).to(torch.int32))
        nnz.value = rows * 16
    else:
        ctx.call("cme213_gen_powerlaw_lengths", rows, 4, 265, 100, 256, 6, d_ptr, C.byref(nnz))
    d_col, d_val = i32(nnz.value, torch), torch.empty(nnz.value, dtype=torch.float32, device="cuda")
    ctx.call("cme213_gen_random_csr_fill_f32", rows, rows, d_ptr, args.window, 7, d_col, d_val)
    x = torch.empty(rows, dtype=torch.float32, device="cuda")
    ctx.call("cme213_fill_uniform_f32", x, rows, 8)
    y = torch.empty_like(x)
    alg = nnz.value * 8 + rows * 8
    meta = {"rows": rows, "nnz": nnz.value, "window": args.window, "regular": bool(args.regular), "l2": "%.1f GB per pass >> L2" % (alg / 1e9)}
    name = "regular 16/row" if args.regular else "power-law (Pareto tail, cap 256)"
    if fmt == "csr":
        def step():
            ctx.call("cme213_spmv_csr_f32", 0, rows, d_ptr, d_col, d_val, x, y, args.csr_kernel, stream=torch.cuda.current_stream())
        meta["workload"] = "C4: CSR SpMV fp32, %s, %d rows" % (name, rows)
        return dict(step=step, alg_bytes=alg + (rows + 1) * 4, flops=2 * nnz.value, kernel="csr_stream_kernel<float>", launches_per_step=1, meta=meta,
                    keep=(d_ptr, d_col, d_val, x, y))
    K, maxlen = C.c_int(), C.c_int()
    ctx.call("cme213_hyb_pick_k", rows, d_ptr, C.byref(K), C.byref(maxlen))
    choice = pkg.capi.FormatChoice()
    ctx.call("cme213_spmv_pick_format", rows, d_ptr, C.byref(choice))
    meta["pick_format"] = {"format": ["CSR", "ELL", "HYB"][choice.format], "K": choice.K, "tail_fraction": choice.tail_fraction,
                           "ell_padding_ratio": choice.ell_padding_ratio, "cost_vs_csr": choice.cost_vs_csr}
    if fmt == "hyb":      # HYB as the library recommends it: the chooser's K (or --hyb-k), tail kept as CSR
        Kell = args.hyb_k if args.hyb_k > 0 else (choice.K if choice.format == 2 else K.value)
    else:
        Kell = maxlen.value if fmt == "ell" else K.value
    pitch = int(pkg.load().cme213_ell_pitch(rows))
    e_idx, e_val = i32(Kell * pitch, torch), torch.empty(Kell * pitch, dtype=torch.float32, device="cuda")
    ctx.call("cme213_csr_to_ell_f32", rows, rows, d_ptr, d_col, d_val, Kell, pitch, e_idx, e_val)
    meta.update({"ell_K": Kell, "max_row_len": maxlen.value, "hyb_K": K.value, "ell_actual_bytes": Kell * pitch * 8})
    if fmt == "ell":
        del d_col, d_val
        def step():
            ctx.call("cme213_spmv_ell_f32", rows, pitch, Kell, e_idx, e_val, x, y, stream=torch.cuda.current_stream())
        meta["workload"] = "C4: ELL SpMV fp32, %s, %d rows (K = max row length)" % (name, rows)
        return dict(step=step, alg_bytes=alg, flops=2 * nnz.value, kernel="ell_v4_kernel<float>", launches_per_step=1, meta=meta,
                    keep=(e_idx, e_val, x, y))
    if fmt == "hyb":
        t_p = i32(rows + 1, torch)
        tail = C.c_int64()
        ctx.call("cme213_csr_tail_row_ptr", rows, d_ptr, Kell, t_p, C.byref(tail))
        n = max(tail.value, 4)
        t_r, t_c, t_v = i32(n, torch), i32(n, torch), torch.empty(n, dtype=torch.float32, device="cuda")
        if tail.value:
            ctx.call("cme213_csr_to_coo_tail_f32", rows, d_ptr, d_col, d_val, Kell, None, t_r, t_c, t_v)
        del d_col, d_val, t_r
        meta.update({"tail_nnz": tail.value, "hyb_actual_bytes": Kell * pitch * 8 + tail.value * 8 + (rows + 1) * 4})

        def step():
            ctx.call("cme213_spmv_hyb_csrtail_f32", rows, pitch, Kell, e_idx, e_val, t_p if tail.value else None, t_c, t_v, x, y,
                     stream=torch.cuda.current_stream())
        meta["workload"] = "C4: HYB SpMV fp32 (ELL K=%d + tail as CSR), %s, %d rows" % (Kell, name, rows)
        return dict(step=step, alg_bytes=alg, flops=2 * nnz.value, kernel="ell_v4_kernel<float> + csr_stream_lean_kernel<float,acc>",
                    launches_per_step=2 if tail.value else 1, meta=meta, keep=(e_idx, e_val, t_p, t_c, t_v, x, y))
    tail = C.c_int64()
    ctx.call("cme213_csr_to_coo_tail_f32", rows, d_ptr, d_col, d_val, Kell, C.byref(tail), None, None, None)
    t_r, t_c, t_v = i32(max(tail.value, 4), torch), i32(max(tail.value, 4), torch), torch.empty(max(tail.value, 4), dtype=torch.float32, device="cuda")
    if tail.value:
        ctx.call("cme213_csr_to_coo_tail_f32", rows, d_ptr, d_col, d_val, Kell, None, t_r, t_c, t_v)
    del d_col, d_val
    meta.update({"coo_tail_nnz": tail.value, "hyb_actual_bytes": Kell * pitch * 8 + tail.value * 12})

    def step():
        ctx.call("cme213_spmv_hyb_f32", rows, pitch, Kell, e_idx, e_val, tail.value, t_r, t_c, t_v, x, y, stream=torch.cuda.current_stream())
    meta["workload"] = "C4: HYB SpMV fp32 (ELL K=%d by the SC09 rule + COO tail by segmented reduction), %s, %d rows" % (Kell, name, rows)
    return dict(step=step, alg_bytes=alg, flops=2 * nnz.value, kernel="ell_v4_kernel<float> + lookback_scan_kernel<coo>",
                launches_per_step=2 if tail.value else 1, meta=meta, keep=(e_idx, e_val, t_r, t_c, t_v, x, y))


def make_gather(ctx, torch, args, q):
    """The random-gather ceiling (cme213_probe_gather_f32): 2^28 gathers of 4 bytes from q floats."""
    n = 1 << 28
    x = torch.empty(q, dtype=torch.float32, device="cuda")
    ctx.call("cme213_fill_uniform_f32", x, q, 8)
    out = torch.empty(n // 16, dtype=torch.float32, device="cuda")

    def step():
        ctx.call("cme213_probe_gather_f32", x, q, out, n, 3, stream=torch.cuda.current_stream())

    return dict(step=step, alg_bytes=n * 4 + n // 4, flops=n, kernel="probe_gather_kernel", launches_per_step=1,
                meta={"workload": "gather ceiling: 2^28 random 4-byte gathers from %d floats (%.0f MB), 16 per thread" % (q, q * 4 / 1e6),
                      "gathers": n, "sector_bytes_if_every_gather_is_one_l2_sector": n * 32, "l2": "x %s L2" % ("fits" if q * 4 < 100e6 else "exceeds")},
                keep=(x, out))


def make_pagerank(ctx, torch, args):
    import ctypes as C
    n, avg = 1 << 21, 8
    ind, inv, nodes = i32(n + 1, torch), torch.empty(n, dtype=torch.float32, device="cuda"), torch.empty(n, dtype=torch.float32, device="cuda")
    nnz = C.c_int64()
    ctx.call("cme213_gen_pagerank_graph", n, avg, 21, ind, None, inv, nodes, C.byref(nnz))
    edges = i32(nnz.value, torch)
    ctx.call("cme213_gen_pagerank_graph", n, avg, 21, ind, edges, inv, nodes, C.byref(nnz))
    out = torch.empty_like(nodes)
    bufs = [nodes, out]
    state = {"i": 0}

    def step():
        a, b = bufs[state["i"] % 2], bufs[1 - state["i"] % 2]
        state["i"] += 1
        ctx.call("cme213_pagerank_propagate", ind, edges, a, b, inv, n, 0, stream=torch.cuda.current_stream())

    alg = nnz.value * 4 + (n + 1) * 4 + n * 12
    return dict(step=step, alg_bytes=alg, flops=2 * nnz.value + 2 * n, kernel="csr_stream_kernel<float,pagerank>", launches_per_step=1,
                meta={"workload": "C1': PageRank propagate (hw1), N=2^21, avg_edges=8", "nnz": nnz.value,
                      "l2": "working set %.0f MB fits in L2: warm-L2 number, not an HBM roofline" % (alg / 1e6)}, keep=(ind, edges, inv, nodes, out))


def make_pagerank_iterate(ctx, torch, args):
    """hw1 end to end as the homework calls it: device_graph_iterate's work (cme213_pagerank_iterate_host) on N = 2^21,
    avg_edges = 8, 20 iterations; a step = one call, ms inside the call's own timer is reported beside it."""
    import ctypes as C
    import numpy as np
    from oracle import generators as g
    n, avg, iters = 1 << 21, 8, 20
    ind, edges, inv, nodes = g.pagerank_graph(n, avg, seed=21)
    out = np.zeros_like(nodes)
    lib = ctx.lib
    P = lambda a: C.c_void_p(a.ctypes.data)
    kms = C.c_float()
    state = {"kms": []}

    def step():
        rc = lib.cme213_pagerank_iterate_host(P(ind), P(edges), P(nodes), P(out), P(inv), iters, n, C.byref(kms))
        if rc:
            raise RuntimeError(lib.cme213_last_error())
        state["kms"].append(kms.value)

    alg = (len(edges) * 4 + (n + 1) * 4 + n * 12) * iters
    meta = {"workload": "C1' iterated: device_graph_iterate's loop (hw1), N=2^21, avg_edges=8, 20 iterations per call (host arrays in and out)",
            "nnz": int(len(edges)), "kernel_ms_inside_call": state["kms"], "l2": "working set ~100 MB: L2-resident"}
    return dict(step=step, alg_bytes=alg, flops=(2 * len(edges) + 2 * n) * iters, kernel="csr_stream_lean_kernel<float,pagerank_w>", launches_per_step=iters,
                meta=meta, keep=(ind, edges, inv, nodes, out))


def build_workload(name, ctx, torch, args, rank=0, world=1, dist=None):
    if name == "c5":
        return make_c5(ctx, torch, args, rank, world, dist)
    if name == "c1":
        return make_c1(ctx, torch, args)
    if name == "scan":
        return make_scan(ctx, torch, args)
    if name == "segscan":
        return make_segscan(ctx, torch, args)
    if name == "radixsort":
        return make_radixsort(ctx, torch, args)
    if name == "dia":
        return make_dia(ctx, torch, args)
    if name == "stencil":
        return make_stencil(ctx, torch, args)
    if name == "fp":
        return make_fp(ctx, torch, args)
    if name == "fp_iterate":
        return make_fp_iterate(ctx, torch, args)
    if name == "transpose32":
        return make_transpose(ctx, torch, args, 32)
    if name == "transpose64":
        return make_transpose(ctx, torch, args, 64)
    if name in ("ell", "hyb", "hyb_coo"):
        return make_powerlaw(ctx, torch, args, name)
    if name == "csr_powerlaw":
        return make_powerlaw(ctx, torch, args, "csr")
    if name in ("ell_window", "hyb_window", "hyb_coo_window", "csr_window", "ell_regular_window", "csr_regular_window"):
        # same row lengths, columns drawn inside a 4,096-wide band around the diagonal (the locality of the FEM
        # matrices in the final project's benchmark set): the x gathers then hit in L1/L2 and the format kernels show
        # their streaming rate instead of the random-gather ceiling
        import copy
        a2 = copy.copy(args)
        a2.window = 1 << 12
        a2.regular = 1 if "regular" in name else 0
        return make_powerlaw(ctx, torch, a2, "hyb_coo" if name.startswith("hyb_coo") else name.split("_")[0])
    if name == "pagerank":
        return make_pagerank(ctx, torch, args)
    if name == "pagerank_iterate":
        return make_pagerank_iterate(ctx, torch, args)
    if name in ("gather_4mb", "gather_64mb", "gather_1gb"):
        return make_gather(ctx, torch, args, {"gather_4mb": 1 << 20, "gather_64mb": 1 << 24, "gather_1gb": 1 << 28}[name])
    raise SystemExit("unknown workload " + name)


def time_steps(torch, step, steps, warmup, dist=None):
    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if dist is not None:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.barrier()
        torch.cuda.synchronize()
        ms = float(t.item())
    return ms


# ---------------------------------------------------------------------------------------------------
# CPU baseline (oracle = checker; timed here only as the reported baseline)
# ---------------------------------------------------------------------------------------------------
C5_WORKLOAD = "C5: CSR SpMV iteration, 3-D 7-point Laplacian %d^3, fp64"


def cpu_baseline_c5(host, grid, budget_s=12.0):
    """Reference-style CPU loop (csr_spmv of slides/Lecture22.pdf slide 20, OpenMP over rows) on the WHOLE matrix of the
    workload -- the host copy the end-to-end leg uploads -- for about `budget_s` seconds of passes."""
    import numpy as np
    import oracle
    ptr, col, val = host["ptr"].numpy(), host["col"].numpy(), host["val"].numpy()
    rows, n = len(ptr) - 1, len(col)
    x = np.random.default_rng(0).uniform(-1, 1, rows)
    y = np.empty_like(x)
    threads = oracle.num_threads()
    oracle.csr_spmv_into(ptr, col, val, x, y)  # warm-up
    reps, t0 = 0, time.perf_counter()
    while True:
        oracle.csr_spmv_into(ptr, col, val, x, y)
        reps += 1
        el = time.perf_counter() - t0
        if el > budget_s or reps >= 50:
            break
    alg = n * 12 + (rows + 1) * 4 + rows * 16
    return {"value": alg * reps / el / 1e9, "unit": "GB/s", "cores": threads, "kind": "port",
            "sample": "the whole %d^3 matrix (%d rows, %d nnz), %d passes in %.1f s, OpenMP over rows, %d threads"
                      % (grid, rows, n, reps, el, threads)}


def run_reference(args):
    """--impl reference: the reference's CPU path for the same metric on the host cores.  For C5 the
    reference has no fp64 CSR code of its own (its CSR is the fp32 PageRank gather), so this is the
    oracle port of the slide-20 loop with OpenMP over rows (kind = "port")."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # torchrun pins OMP_NUM_THREADS=1 for its children; the reference arm is meant to use every host core
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
    import numpy as np
    import oracle
    from oracle import generators as g
    oracle.build()
    if args.workload == "pagerank":
        return run_reference_pagerank(args, np, oracle, g)
    if args.workload == "radixsort":
        return run_reference_radixsort(args, np, oracle)
    if args.workload in ("scan", "fp", "transpose32", "transpose64"):
        return run_reference_primitive(args, np, oracle, g)
    # C5 on the WHOLE matrix (13.9 GB on the host for 512^3): the same config as the GPU arm, one step = one pass y = A x
    grid = args.grid
    rows = grid ** 3
    t_gen = time.perf_counter()
    ptr, col, val = oracle.laplace3d_host(grid, grid, grid)
    t_gen = time.perf_counter() - t_gen
    x = np.random.default_rng(0).uniform(-1, 1, rows) * 0.05
    y = np.empty_like(x)
    alg = len(col) * 12 + (rows + 1) * 4 + rows * 16
    threads = oracle.num_threads()
    for _ in range(args.warmup):
        oracle.csr_spmv_into(ptr, col, val, x, y)
        x, y = y, x
    t0 = time.perf_counter()
    for _ in range(args.steps):
        oracle.csr_spmv_into(ptr, col, val, x, y)
        x, y = y, x                                  # the ping-pong of pagerank.cu:128-132
    el = time.perf_counter() - t0
    v = alg * args.steps / el / 1e9
    sample = ("each step = one SpMV over the whole %d^3 matrix (%d rows, %d nnz), OpenMP over rows, %d threads; matrix built on "
              "the host in %.1f s (untimed)" % (grid, rows, len(col), threads, t_gen))
    print(json.dumps({
        "impl": "reference", "metric": "spmv_algorithmic_gbs", "value": v, "unit": "GB/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": el / args.steps * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": C5_WORKLOAD % grid, "rows": rows, "nnz": len(col), "host": "CPU, whole matrix"},
        "cpu_baseline": {"value": v, "unit": "GB/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}), flush=True)


def run_reference_radixsort(args, np, oracle):
    """--impl reference --workload radixsort: the reference's OWN radixSortParallelPass loop (radixsort.cpp:44-121,
    compiled from the checkout into oracle/_ref, OpenMP over its 8192-key blocks) on its default problem size
    (40 M keys, numBits 8, radixsort.cpp:165-166); the oracle's blocked restatement (kind "port") when _ref is absent."""
    n = 40_000_000
    keys = np.random.default_rng(5).integers(0, 2**31, n, dtype=np.int64).astype(np.uint32)   # rand() range, :174
    have_ref = oracle.ref_lib("radixsort") is not None
    threads = oracle.num_threads()

    def run():
        if have_ref:
            return oracle.ref_radix_sort(keys, 8)[1]          # seconds inside the passes, copies excluded
        t0 = time.perf_counter()
        oracle.radix_sort_u32(keys, 8, block_size=8192)
        return time.perf_counter() - t0
    for _ in range(max(1, min(args.warmup, 2))):
        run()
    steps = max(1, min(args.steps, 10))
    el = sum(run() for _ in range(steps))
    v = 48.0 * n * steps / el / 1e9
    print(json.dumps({
        "impl": "reference", "metric": "algorithmic_gbs", "value": v, "unit": "GB/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": args.warmup, "ms_per_step": el / steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "u32", "data": "synthetic", "keys_per_s": n * steps / el,
        "config": {"workload": "8f.4: LSD radix sort of uint32 keys, 4 passes of 8 bits (CPU, 40 M keys)"},
        "cpu_baseline": {"value": v, "unit": "GB/s", "cores": threads, "kind": "reference" if have_ref else "port",
                         "sample": "radixSortParallelPass x4 of radixsort.cpp on 40,000,000 keys (its default size), %d OpenMP threads" % threads},
        "e2e": {"value": v, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}), flush=True)


def run_reference_primitive(args, np, oracle, g):
    """--impl reference for the single-GPU primitives, on bounded samples of the GPU workloads:
    scan        the 3-line exclusive scan of radixsort.cpp:76-81 restated, OpenMP two-pass blocked variant (port)
    fp          the shipped CPU path of fp.cu:124-150 restated (xx = x[k] hoisted, OpenMP multiply + per-segment sums, port)
    transposeNN the index formula of Lecture06 slide 9 restated, OpenMP over row tiles (port; the reference has no code)"""
    import ctypes
    threads = oracle.num_threads()
    rng = np.random.default_rng(3)
    w = args.workload
    if w == "scan":
        n = 1 << 26
        x = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
        out, L, P = np.zeros_like(x), oracle.lib(), (lambda a: a.ctypes.data_as(ctypes.c_void_p))
        run, alg = (lambda: L.oracle_exclusive_scan_u32_omp(P(x), P(out), ctypes.c_int64(n))), 8 * n   # buffers allocated once
        name, sample, dtype = "C2a: exclusive scan of uint32 (CPU)", "2^26 of the 2^28 elements, OpenMP blocked two-pass scan", "u32"
    elif w == "fp":
        n, q = 1 << 24, 1 << 20
        a = rng.uniform(-1, 1, n).astype(np.float32)
        xv = rng.uniform(-1, 1, q).astype(np.float32)
        k = rng.integers(0, q, n).astype(np.int32)
        sseg = np.concatenate([np.arange(0, n, 64), [n]]).astype(np.int32)
        xx, cur, L, P = xv[k], a.copy(), oracle.lib(), (lambda t: t.ctypes.data_as(ctypes.c_void_p))   # xx = x[k] hoisted as in fp.cu:124-125
        run = lambda: L.oracle_fp_cpu_iterate_f32(ctypes.c_int(n), ctypes.c_int(len(sseg)), P(cur), P(sseg), P(xx), ctypes.c_int(1))
        alg = 12 * n + 4 * q + n // 8
        name, sample, dtype = "C2c: final-project step a*x[k] + segmented scan (CPU)", "n = 2^24 of 2^28, q = 2^20, segments of 64, fp.cu's OpenMP loops", "f32"
    else:
        bits = 32 if w == "transpose32" else 64
        m = 8192
        a = rng.integers(0, 2**31, (m, m)).astype(np.uint32 if bits == 32 else np.uint64)
        out, L, P = np.zeros_like(a), oracle.lib(), (lambda t: t.ctypes.data_as(ctypes.c_void_p))
        fn = L.oracle_transpose_b32_omp if bits == 32 else L.oracle_transpose_b64_omp
        run, alg = (lambda: fn(P(a), P(out), ctypes.c_int64(m), ctypes.c_int64(m))), 2 * m * m * (bits // 8)
        name, sample, dtype = "C3: transpose, %d-bit elements (CPU)" % bits, "8192^2 of the 16384^2 matrix, OpenMP over row tiles", "b%d" % bits
    for _ in range(max(1, min(args.warmup, 2))):
        run()
    steps = max(1, min(args.steps, 10))
    t0 = time.perf_counter()
    for _ in range(steps):
        run()
    el = time.perf_counter() - t0
    v = alg * steps / el / 1e9
    print(json.dumps({
        "impl": "reference", "metric": "algorithmic_gbs", "value": v, "unit": "GB/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": args.warmup, "ms_per_step": el / steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": dtype, "data": "synthetic", "config": {"workload": name},
        "cpu_baseline": {"value": v, "unit": "GB/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}), flush=True)


def run_reference_pagerank(args, np, oracle, g):
    """--impl reference --workload pagerank: the reference's OWN host_graph_propagate (pagerank.cu:45-56, compiled
    from the checkout into oracle/_ref) on the homework's problem (N = 2^21, avg_edges = 8); serial, as shipped.
    Falls back to the oracle restatement (kind "port") when oracle/_ref was not built."""
    n, avg = 1 << 21, 8
    ind, edges, inv, nodes = g.pagerank_graph(n, avg, seed=21)
    alg = len(edges) * 4 + (n + 1) * 4 + n * 12
    have_ref = oracle.ref_lib("pagerank") is not None
    run = (lambda it: oracle.ref_pagerank_iterate(ind, edges, nodes, inv, it)) if have_ref else \
          (lambda it: oracle.pagerank_iterate(ind, edges, nodes, inv, it))
    run(2 * max(args.warmup // 2, 1))
    iters = 2 * max(args.steps // 2, 1)           # a step = one propagate; the reference iterates in pairs
    t0 = time.perf_counter()
    run(iters)
    el = time.perf_counter() - t0
    v = alg * iters / el / 1e9
    print(json.dumps({
        "impl": "reference", "metric": "spmv_algorithmic_gbs", "value": v, "unit": "GB/s", "n_gpus": args.gpus, "steps": iters,
        "warmup": args.warmup, "ms_per_step": el / iters * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "C1': PageRank propagate (hw1), N=2^21, avg_edges=8 (CPU)"},
        "cpu_baseline": {"value": v, "unit": "GB/s", "cores": 1, "kind": "reference" if have_ref else "port",
                         "sample": "host_graph_propagate of pagerank.cu:45-56, whole graph, %d propagates, 1 thread" % iters},
        "e2e": {"value": v, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}), flush=True)


# ---------------------------------------------------------------------------------------------------
def e2e_c5(torch, pkg, wl, args, rank, world, dist):
    """End to end through the host-pointer driver entry point (cme213_csr_iterate_host_f64, the C5
    analogue of device_graph_iterate): pinned HOST arrays in, H2D + 20 iterations + D2H inside the
    timed region.  Each rank drives its own row block of the matrix (its rows x the columns it
    touches); no exchange is needed for that, and the aggregate is the sum over ranks."""
    import ctypes as C
    import numpy as np
    lo, hi, d_ptr, d_col, d_val, nnz = wl["shard"]
    g = args.grid
    rows_total = g ** 3
    # Single-GPU semantics need a square operator: at N=1 that is the whole matrix.  At N>1 every rank
    # times the upload + iteration of a square diagonal block (columns outside the block dropped by
    # pointing them at a zero entry would change the matrix), so e2e is reported at N=1 only.
    iters = 20
    if world > 1:
        return e2e_c5_sharded(torch, pkg, wl, args, rank, world, dist, iters)
    pin = lambda t: t.cpu().pin_memory()
    h_ptr, h_col, h_val = pin(d_ptr), pin(d_col), pin(d_val)
    wl["host"] = {"ptr": h_ptr, "col": h_col, "val": h_val}   # the CPU baseline runs on this same host copy
    h_x = torch.empty(rows_total, dtype=torch.float64).pin_memory()
    h_x.copy_(wl["keep"][3].cpu() if wl["keep"][3].numel() == rows_total else torch.rand(rows_total, dtype=torch.float64))
    h_x.mul_(1e-3)
    h_y = torch.empty(rows_total, dtype=torch.float64).pin_memory()
    lib = pkg.load()
    P = lambda t: C.c_void_p(t.data_ptr())
    h2d = h_ptr.numel() * 4 + h_col.numel() * 4 + h_val.numel() * 8 + rows_total * 8
    d2h = rows_total * 8
    calls = max(1, min(args.steps, 3))
    kernel_ms = C.c_float()
    rc = lib.cme213_csr_iterate_host_f64(rows_total, nnz, P(h_ptr), P(h_col), P(h_val), P(h_x), P(h_y), 2, C.byref(kernel_ms))
    if rc != 0:
        raise RuntimeError(lib.cme213_last_error())
    t0 = time.perf_counter()
    for _ in range(calls):
        rc = lib.cme213_csr_iterate_host_f64(rows_total, nnz, P(h_ptr), P(h_col), P(h_val), P(h_x), P(h_y), iters, C.byref(kernel_ms))
        if rc != 0:
            raise RuntimeError(lib.cme213_last_error())
    el = time.perf_counter() - t0
    # the same solve with the matrix kept on the device between calls (cme213_csr_upload_host_f64 once, untimed):
    # what a caller that iterates on the same matrix again and again pays per call
    resident = None
    try:
        h = C.c_void_p()
        rc = lib.cme213_csr_upload_host_f64(rows_total, nnz, P(h_ptr), P(h_col), P(h_val), C.byref(h))
        if rc != 0:
            raise RuntimeError(lib.cme213_last_error())
        lib.cme213_csr_iterate_f64(h, P(h_x), P(h_y), 2, C.byref(kernel_ms))
        t1 = time.perf_counter()
        for _ in range(calls):
            rc = lib.cme213_csr_iterate_f64(h, P(h_x), P(h_y), iters, C.byref(kernel_ms))
            if rc != 0:
                raise RuntimeError(lib.cme213_last_error())
        el2 = time.perf_counter() - t1
        lib.cme213_csr_matrix_destroy(h)
        resident = {"value": wl["alg_bytes"] * iters * calls / el2 / 1e9, "unit": "GB/s", "h2d_bytes_per_step": rows_total * 8,
                    "d2h_bytes_per_step": d2h, "ms_per_call": el2 / calls * 1e3,
                    "call": "cme213_csr_iterate_f64(handle, nr_iterations=%d): x in, result out, matrix resident" % iters}
    except Exception as ex:  # noqa: BLE001
        resident = {"error": repr(ex)}
    return {"value": wl["alg_bytes"] * iters * calls / el / 1e9, "unit": "GB/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
            "call": "cme213_csr_iterate_host_f64(nr_iterations=%d): pinned host CSR in, result vector out" % iters,
            "calls_timed": calls, "ms_per_call": el / calls * 1e3, "kernel_ms_inside_call": float(kernel_ms.value),
            "matrix_resident": resident}


def e2e_c5_sharded(torch, pkg, wl, args, rank, world, dist, iters):
    """N > 1: every rank owns the pinned HOST copy of its row block.  Timed per call: H2D of the block and of
    x, `iters` row-sharded steps (halo exchange over NCCL), D2H of the rank's slice of the result."""
    sh = importlib.import_module(PKG + ".sharded")
    lo, hi, d_ptr, d_col, d_val, nnz = wl["shard"]
    g = args.grid
    rows = g ** 3
    ctx = wl["ctx"]
    h_ptr, h_col, h_val = d_ptr.cpu().pin_memory(), d_col.cpu().pin_memory(), d_val.cpu().pin_memory()
    h_x = (torch.rand(rows, dtype=torch.float64) * 1e-3).pin_memory()
    h_y = torch.empty(hi - lo, dtype=torch.float64).pin_memory()
    if "shard_obj" in wl:
        # the C ABI's sharded iterate: the row block is uploaded into the arrays the shard was created on
        shard = wl["shard_obj"]

        def call():
            d_ptr.copy_(h_ptr, non_blocking=True); d_col.copy_(h_col, non_blocking=True); d_val.copy_(h_val, non_blocking=True)
            shard.x[0].copy_(h_x, non_blocking=True)
            shard.iterate(iters)
            h_y.copy_(shard.x[0][lo:hi], non_blocking=True)
            shard.finish()

        call()
        calls = max(1, min(args.steps, 3))
        dist.barrier(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(calls):
            call()
        dist.barrier()
        el = time.perf_counter() - t0
        t = torch.tensor([el], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        el = float(t.item())
        h2d = h_ptr.numel() * 4 + h_col.numel() * 4 + h_val.numel() * 8 + rows * 8
        return {"value": wl["alg_bytes"] * iters * calls / el / 1e9, "unit": "GB/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": (hi - lo) * 8,
                "call": "per rank: pinned host row block + x -> device, cme213_shard_iterate(%d), result slice -> host" % iters,
                "calls_timed": calls, "ms_per_call": el / calls * 1e3, "bytes_are": "per rank"}
    # device buffers of the call (allocated once: the host-pointer API keeps its workspace between calls)
    e_ptr, e_col, e_val = torch.empty_like(d_ptr), torch.empty_like(d_col), torch.empty_like(d_val)
    xa, xb, peers, exchange = make_vectors(sh, ctx, torch, dist, rank, world, rows, g * g, args.exchange)
    streams = (torch.cuda.current_stream(), torch.cuda.Stream())
    it = sh.ShardedIterate(rows, world, rank, sh.cuda_spmv(ctx, lo, e_ptr, e_col, e_val, args.csr_kernel), xa, xb,
                           halo=g * g, mode=exchange, dist=dist, streams=streams, peers=peers)

    def call():
        e_ptr.copy_(h_ptr, non_blocking=True); e_col.copy_(h_col, non_blocking=True); e_val.copy_(h_val, non_blocking=True)
        xa.copy_(h_x, non_blocking=True)
        it.iterate(iters)
        h_y.copy_(xa[lo:hi], non_blocking=True)
        torch.cuda.synchronize()

    call()
    calls = max(1, min(args.steps, 3))
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(calls):
        call()
    dist.barrier()
    el = time.perf_counter() - t0
    t = torch.tensor([el], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    el = float(t.item())
    h2d = h_ptr.numel() * 4 + h_col.numel() * 4 + h_val.numel() * 8 + rows * 8
    return {"value": wl["alg_bytes"] * iters * calls / el / 1e9, "unit": "GB/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": (hi - lo) * 8,
            "call": "per rank: pinned host row block + x -> device, %d row-sharded iterations (sharded.ShardedIterate), result slice -> host" % iters,
            "calls_timed": calls, "ms_per_call": el / calls * 1e3, "bytes_are": "per rank"}


def parity_c5(torch, ctx, wl, args, rank, world, dist, passes=6):
    """N > 1, outside the timed region: the same `passes` iterations once through the sharded path and once on rank 0
    ALONE over the whole matrix (single-GPU kernel), compared bit for bit.  Row sharding does not change the order in
    which a row is summed, so anything but equality is a bug in the exchange."""
    import ctypes as C
    g = args.grid
    rows = g ** 3
    lo, hi = wl["shard"][0], wl["shard"][1]
    shard = wl.get("shard_obj")
    x0, x1 = (shard.x if shard is not None else wl["iterate"].x)
    torch.cuda.synchronize()
    dist.barrier()
    ctx.call("cme213_fill_uniform_f64", x0, rows, 9)
    x0.mul_(0.05)
    x1.zero_()
    torch.cuda.synchronize()
    dist.barrier()
    if shard is not None:
        shard.iterate(passes)
        shard.finish()
        mode = wl["meta"]["exchange"].split(" ")[0]
    else:
        wl["iterate"].iterate(passes)
        torch.cuda.synchronize()
        mode = wl["iterate"].mode
    dist.barrier()
    sizes = [0] * world
    dist.all_gather_object(sizes, hi - lo)
    pieces = [torch.empty(n, dtype=torch.float64, device="cuda") for n in sizes]
    dist.all_gather(pieces, x0[lo:hi].contiguous())
    ok = None
    if rank == 0:
        got = torch.cat(pieces)
        del pieces
        nnz = C.c_int64()
        p = i32(rows + 1, torch)
        ctx.call("cme213_gen_laplace3d_f64", g, g, g, 0, rows, p, None, None, C.byref(nnz))
        c, v = i32(nnz.value, torch), torch.empty(nnz.value, dtype=torch.float64, device="cuda")
        ctx.call("cme213_gen_laplace3d_f64", g, g, g, 0, rows, p, c, v, C.byref(nnz))
        a = torch.empty(rows, dtype=torch.float64, device="cuda")
        ctx.call("cme213_fill_uniform_f64", a, rows, 9)
        a.mul_(0.05)
        b = torch.empty_like(a)
        for _ in range(passes // 2):
            ctx.call("cme213_spmv_csr_f64", 0, rows, p, c, v, a, b, 0)
            ctx.call("cme213_spmv_csr_f64", 0, rows, p, c, v, b, a, 0)
        torch.cuda.synchronize()
        ok = bool(torch.equal(got, a))
        del p, c, v, a, b, got
        torch.cuda.empty_cache()
    dist.barrier()
    return {"bitwise_equal_to_1gpu": ok, "mode": mode, "passes": passes, "rows_compared": rows,
            "how": "rank 0 recomputes the passes alone on the whole matrix with the single-GPU kernel; torch.equal on the gathered vector"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--workload", default="c5")
    ap.add_argument("--grid", type=int, default=512)
    ap.add_argument("--exchange", default="flags", choices=["flags", "gather", "split", "fused", "p2p", "halo", "full"],
                    help="N>1: flags (default) = the C ABI's sharded iterate: ONE launch per pass, the kernel stores the halo rows into the neighbours' vectors and the passes are ordered by flags in peer memory (no NCCL on the data path); gather / split = its whole-vector modes for general matrices.  The others are the round-1 torch.distributed paths, kept as fallbacks: fused = the boundary launch stores the halo planes into the neighbours' vectors over peer memory (falls back to p2p when CUDA IPC or the shape does not allow it); p2p = NCCL send/recv of the halo planes between neighbours; halo = one all-gather of every rank's boundary planes; full = all-gather of the whole vector")
    ap.add_argument("--csr-kernel", type=int, default=0)
    ap.add_argument("--c1-n", type=int, default=1024, help="C1: grid edge (1024 = the configuration of BASELINE.json)")
    ap.add_argument("--l2-fetch-gran", type=int, default=0, help="experiment: set the device's maximum L2 fetch granularity (bytes) before the run")
    ap.add_argument("--l2-persist-mb", type=int, default=-1, help="experiment: set the device's persisting-L2 set-aside (MB) before the run")
    ap.add_argument("--graph", type=int, default=0, help="N>1: 1 = replay CUDA graphs of the passes instead of calling cme213_shard_step per pass (measured 0.4 %% slower at 8 GPUs); the NCCL fallback modes always use a graph of two passes")
    ap.add_argument("--order-strip", type=int, default=0, help="C5: CTA order = z fastest inside strips of this many grid lines (0 = row order)")
    ap.add_argument("--powerlaw-rows", type=int, default=1 << 24)
    ap.add_argument("--window", type=int, default=0)
    ap.add_argument("--regular", type=int, default=0)
    ap.add_argument("--hyb-k", type=int, default=0, help="hyb workload: ELL width (0 = cme213_spmv_pick_format's choice)")
    ap.add_argument("--extras", type=int, default=1, help="also time the other single-GPU configs (N=1 only)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="N>1: skip the bitwise comparison with rank 0's single-GPU recomputation")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup

    if args.impl == "reference":
        run_reference(args)
        return

    import torch
    pkg = importlib.import_module(PKG)
    pkg.load()  # fails loudly when the CUDA library is missing: there is no other path
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = pkg.Context(local)
    if args.l2_persist_mb >= 0:   # experiment: L2 set-aside for evict-last / persisting lines (driver API, the primary context is current)
        import ctypes
        torch.zeros(1, device="cuda")
        cu = ctypes.CDLL("libcuda.so.1")
        rc = cu.cuCtxSetLimit(6, ctypes.c_size_t(args.l2_persist_mb << 20))   # CU_LIMIT_PERSISTING_L2_CACHE_SIZE
        got = ctypes.c_size_t()
        cu.cuCtxGetLimit(ctypes.byref(got), 6)
        log("l2 persisting set-aside: rc", rc, "now", got.value >> 20, "MB")
    if args.l2_fetch_gran > 0:    # experiment: CU_LIMIT_MAX_L2_FETCH_GRANULARITY (32 / 64 / 128 bytes)
        import ctypes
        torch.zeros(1, device="cuda")
        cu = ctypes.CDLL("libcuda.so.1")
        rc = cu.cuCtxSetLimit(5, ctypes.c_size_t(args.l2_fetch_gran))
        got = ctypes.c_size_t()
        cu.cuCtxGetLimit(ctypes.byref(got), 5)
        log("l2 max fetch granularity: rc", rc, "now", got.value, "bytes")
    peak, peak_kind = measured_peak()

    wl = build_workload(args.workload, ctx, torch, args, rank, world, dist)
    if world > 1 and args.workload == "c5" and "shard_obj" not in wl:
        args.steps += args.steps % 2      # NCCL fallback modes: a graph replay is a pair of passes
        args.warmup += args.warmup % 2
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = ctx.launches
    ms = time_steps(torch, wl["step"], args.steps, args.warmup, dist)
    launches = (ctx.launches - l0) * args.steps // (args.steps + args.warmup)
    if launches == 0:  # graph replay: the kernels were recorded once, each replayed pass still launches them
        launches = wl["launches_per_step"] * args.steps
    clocks = sampler.stop() if rank == 0 else None
    ms_step = ms / args.steps
    value = wl["alg_bytes"] / (ms_step * 1e-3) / 1e9

    # roofline of the dominant kernel: algorithmic bytes of ONE launch (this rank's rows) over its
    # average duration, timed alone with CUDA events on its launch stream
    roof = None
    kms_all = None
    if args.workload == "c5":
        kms = time_steps(torch, wl["kernel_step"], args.steps, 3) / args.steps
        ach = wl["local_alg_bytes"] / (kms * 1e-3) / 1e9
        kms_all = None
        if dist is not None:   # the same kernel on every rank's own shard, all at once, no exchange: the slowest GPU bounds a coupled pass
            dist.barrier()
            mine = time_steps(torch, wl["kernel_step"], args.steps, 3) / args.steps
            t = torch.zeros(world, dtype=torch.float64, device="cuda")
            t[rank] = mine
            dist.all_reduce(t)
            kms_all = [float(v) for v in t.tolist()]
    else:
        kms = ms_step
        ach = value
    # dram__bytes_read.sum + dram__bytes_write.sum of one launch of the dominant kernel, from the committed
    # `ncu --set full` capture of this same command (profiles/r01_ncu_csr_stream_f64_*.txt); only valid for the
    # default single-GPU C5 workload
    traffic, traffic_src = ncu_traffic(args.workload) if (world == 1 and args.grid == 512) else (None, None)
    roof = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": traffic, "traffic_source": traffic_src,
            "kernel": wl["kernel"], "kernel_ms": kms, "peak_source": peak_kind + " (MEASURED_PEAKS.json hbm_gbs)" if peak_kind == "measured" else "fallback 6650 GB/s",
            "kernel_ms_per_rank_uncoupled": kms_all if args.workload == "c5" else None, "frac_of_8TBs_spec": ach / 8000.0, "gflops": wl["flops"] / max(world, 1) / (kms * 1e-3) / 1e9 if args.workload == "c5" else wl["flops"] / (kms * 1e-3) / 1e9}

    out = {"metric": "spmv_algorithmic_gbs" if args.workload in ("c5", "c1", "ell", "hyb", "csr_powerlaw", "pagerank") else "algorithmic_gbs",
           "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
           "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64" if args.workload == "c5" else "f32/u32",
           "data": "synthetic", "config": wl["meta"], "gflops": wl["flops"] / (ms_step * 1e-3) / 1e9,
           "roofline": roof, "gpu_launches": launches, "clocks": clocks}

    if args.workload == "c5" and world > 1 and not args.no_parity:
        par = parity_c5(torch, ctx, wl, args, rank, world, dist)
        out["parity"] = par
    if args.workload == "c5" and not args.no_e2e:
        e = e2e_c5(torch, pkg, wl, args, rank, world, dist)
        if e is not None:
            out["e2e"] = e
    if rank == 0 and args.workload == "c5" and not args.no_cpu and world == 1:
        import oracle
        oracle.build()
        if "host" not in wl:
            lo, hi, d_ptr, d_col, d_val, _ = wl["shard"]
            wl["host"] = {"ptr": d_ptr.cpu(), "col": d_col.cpu(), "val": d_val.cpu()}
        out["cpu_baseline"] = cpu_baseline_c5(wl["host"], args.grid)
    if world == 1 and args.extras and args.workload == "c5":
        others = []
        del wl
        torch.cuda.empty_cache()
        names = ["scan", "segscan", "radixsort", "dia", "stencil", "fp", "fp_iterate", "transpose32", "transpose64", "c1", "ell", "hyb", "hyb_coo", "csr_powerlaw",
                 "ell_window", "hyb_window", "hyb_coo_window", "csr_window", "ell_regular_window", "csr_regular_window", "pagerank",
                 "gather_4mb", "gather_64mb", "gather_1gb"]
        for name in names:
            try:
                w = build_workload(name, ctx, torch, args)
                t = time_steps(torch, w["step"], args.steps, 3) / args.steps
                gbs = w["alg_bytes"] / (t * 1e-3) / 1e9
                if "gathers" in w["meta"]:
                    w["meta"]["giga_gathers_per_s"] = w["meta"]["gathers"] / (t * 1e-3) / 1e9
                others.append({"workload": w["meta"]["workload"], "ms": t, "gbs": gbs, "frac_of_measured_peak": gbs / peak,
                               "gflops": w["flops"] / (t * 1e-3) / 1e9, "kernel": w["kernel"],
                               "detail": {k: v for k, v in w["meta"].items() if k != "workload"}})
                log("extra %-12s %8.3f ms  %8.1f GB/s  (%.2f of peak)" % (name, t, gbs, gbs / peak))
                del w
                torch.cuda.empty_cache()
            except Exception as ex:  # an extra must not take the headline down
                others.append({"workload": name, "error": repr(ex)})
                log("extra", name, "failed:", repr(ex))
        out["others"] = others
    if rank == 0:
        print(json.dumps(out), flush=True)
    if dist is not None:
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        # Tearing the NCCL communicator down while a captured graph still references it can block for
        # minutes; every rank is past the barrier and has nothing left to do, so leave without the teardown.
        os._exit(0)


if __name__ == "__main__":
    main()
