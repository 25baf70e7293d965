"""Where does a row-sharded pass spend its time?  Run under torchrun like scripts/verify_sharded.py.
Times, on every rank with CUDA events: boundary kernels + pack, the all-gather + unpack alone, the interior
kernel alone, the whole-shard kernel in one launch, and the full overlapped step."""
import ctypes as C
import importlib
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "2012-04_stanford_cme213_b200"


def timed(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    g = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = importlib.import_module(PKG)
    sh = importlib.import_module(PKG + ".sharded")
    ctx = pkg.Context(local)
    rows = g ** 3
    lo, hi = sh.row_partition(rows, world)[rank]
    nnz = C.c_int64()
    d_ptr = torch.empty(hi - lo + 1, dtype=torch.int32, device="cuda")
    ctx.call("cme213_gen_laplace3d_f64", g, g, g, lo, hi, d_ptr, None, None, C.byref(nnz))
    d_col = torch.empty(nnz.value, dtype=torch.int32, device="cuda")
    d_val = torch.empty(nnz.value, dtype=torch.float64, device="cuda")
    ctx.call("cme213_gen_laplace3d_f64", g, g, g, lo, hi, d_ptr, d_col, d_val, C.byref(nnz))
    xa = torch.rand(rows, dtype=torch.float64, device="cuda") * 0.05
    xb = torch.zeros_like(xa)
    streams = (torch.cuda.current_stream(), torch.cuda.Stream())
    it = sh.ShardedIterate(rows, world, rank, sh.cuda_spmv(ctx, lo, d_ptr, d_col, d_val), xa, xb, halo=g * g, mode="halo", dist=dist, streams=streams)
    h = g * g
    spmv = it.spmv
    res = {}
    res["whole shard, one launch"] = timed(lambda: spmv(lo, hi, xa, xb))
    res["interior only"] = timed(lambda: spmv(lo + h, hi - h, xa, xb))

    def boundary():
        spmv(lo, lo + h, xa, xb); spmv(hi - h, hi, xa, xb)
        it.send[:h].copy_(xb[lo:lo + h]); it.send[h:].copy_(xb[hi - h:hi])
    res["2 boundary launches + pack"] = timed(boundary)

    def exchange():
        dist.all_gather_into_tensor(it.recv, it.send)
        if rank > 0:
            xb[lo - h:lo].copy_(it.recv[(rank - 1) * 2 * h + h:(rank - 1) * 2 * h + 2 * h])
        if rank < world - 1:
            xb[hi:hi + h].copy_(it.recv[(rank + 1) * 2 * h:(rank + 1) * 2 * h + h])
    res["all-gather + unpack"] = timed(exchange)
    res["boundary, one launch"] = timed(lambda: spmv.two_ranges(lo, hi, h, xa, xb))
    for mode in ("halo", "p2p"):
        itm = sh.ShardedIterate(rows, world, rank, spmv, xa, xb, halo=h, mode=mode, dist=dist, streams=streams)
        state = {"s": 0}

        def step():
            itm.step(state["s"], 1 - state["s"]); state["s"] ^= 1
        res["%s step (eager)" % mode] = timed(step)
        replay = itm.capture_pair()
        if replay is not None:
            res["%s step (graph, per pass)" % mode] = timed(replay) / 2
    for r in range(world):
        dist.barrier()
        if r == rank and rank in (0, world // 2):
            print("rank %d/%d grid %d^3 rows/rank %d:" % (rank, world, g, hi - lo), "; ".join("%s %.3f ms" % kv for kv in res.items()), flush=True)
    dist.barrier()
    torch.cuda.synchronize()
    os._exit(0)


if __name__ == "__main__":
    main()
