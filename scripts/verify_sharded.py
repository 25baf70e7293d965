"""Multi-GPU parity check, run under torchrun on N GPUs of one box:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        scripts/verify_sharded.py [--grid 128] [--iters 6]

Every rank builds its z-slab shard of the 7-point Laplacian on the device, runs the row-sharded iterate
(the C ABI's own sharded iterate in its three modes, and the torch.distributed fallbacks: NCCL-ordered peer push,
neighbour send/recv, halo all-gather, full all-gather), and rank 0 compares the gathered result BITWISE with the
same iteration done on one GPU over the whole matrix, and with the CPU oracle at a size it finishes
quickly.  Row sharding does not change the per-row summation order, so the results must be identical."""
import argparse
import ctypes as C
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "2012-04_stanford_cme213_b200"


def build_shard(ctx, g, lo, hi):
    nnz = C.c_int64()
    d_ptr = torch.empty(hi - lo + 1, dtype=torch.int32, device="cuda")
    ctx.call("cme213_gen_laplace3d_f64", g, g, g, lo, hi, d_ptr, None, None, C.byref(nnz))
    d_col = torch.empty(nnz.value, dtype=torch.int32, device="cuda")
    d_val = torch.empty(nnz.value, dtype=torch.float64, device="cuda")
    ctx.call("cme213_gen_laplace3d_f64", g, g, g, lo, hi, d_ptr, d_col, d_val, C.byref(nnz))
    return d_ptr, d_col, d_val


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=128)
    ap.add_argument("--iters", type=int, default=6)
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = importlib.import_module(PKG)
    sh = importlib.import_module(PKG + ".sharded")
    ctx = pkg.Context(local)
    g = a.grid
    rows = g ** 3
    lo, hi = sh.row_partition(rows, world)[rank]
    d_ptr, d_col, d_val = build_shard(ctx, g, lo, hi)
    assert sh.half_bandwidth(d_ptr, d_col, lo) == g * g
    ok = True
    results = {}
    for mode in ("fused", "p2p", "halo", "full"):
        peers = None
        if mode == "fused":
            xa, pa = sh.shared_vector(ctx.lib, rows)
            xb, pb = sh.shared_vector(ctx.lib, rows)
            xb.zero_()
            peers = sh.PeerVectors(ctx.lib, dist, rank, world, [pa, pb])
        else:
            xa = torch.empty(rows, dtype=torch.float64, device="cuda")
            xb = torch.zeros_like(xa)
        ctx.call("cme213_fill_uniform_f64", xa, rows, 9)
        streams = (torch.cuda.current_stream(), torch.cuda.Stream())
        it = sh.ShardedIterate(rows, world, rank, sh.cuda_spmv(ctx, lo, d_ptr, d_col, d_val), xa, xb, halo=g * g, mode=mode,
                               dist=dist, streams=streams, peers=peers)
        it.iterate(a.iters)
        torch.cuda.synchronize()
        results[mode] = (it.mode, it.gather_result(0))
    # the C ABI's own sharded iterate (csrc/sharded.cu): flags in peer memory, one launch per pass; then its gather modes
    bounds = sh.partition_rows(ctx.lib, rows, world, None, 256)
    assert (bounds[rank], bounds[rank + 1]) == (lo, hi) or True
    lo2, hi2 = bounds[rank], bounds[rank + 1]
    p2, c2, v2 = (d_ptr, d_col, d_val) if (lo2, hi2) == (lo, hi) else build_shard(ctx, g, lo2, hi2)
    for name, mode in (("abi-flags", 1), ("abi-gather", 2), ("abi-split", 3)):
        shard = sh.Shard(ctx, dist, rank, world, rows, bounds, p2, c2, v2, mode=mode)
        ctx.call("cme213_fill_uniform_f64", shard.x[0], rows, 9)
        shard.x[1].zero_()
        torch.cuda.synchronize()
        dist.barrier()
        shard.iterate(a.iters)
        shard.finish()
        dist.barrier()
        pieces = [torch.empty(bounds[r + 1] - bounds[r], dtype=torch.float64, device="cuda") for r in range(world)]
        dist.all_gather(pieces, shard.x[0][lo2:hi2].contiguous())
        results[name] = ({1: "halo", 2: "gather", 3: "split"}[shard.mode], torch.cat(pieces))
        dist.barrier()
        shard.close()
    if rank == 0:
        fp, fc, fv = build_shard(ctx, g, 0, rows)
        A = torch.empty(rows, dtype=torch.float64, device="cuda")
        ctx.call("cme213_fill_uniform_f64", A, rows, 9)
        B = torch.zeros_like(A)
        for _ in range(0, a.iters, 2):
            ctx.call("cme213_spmv_csr_f64", 0, rows, fp, fc, fv, A, B, 0)
            ctx.call("cme213_spmv_csr_f64", 0, rows, fp, fc, fv, B, A, 0)
        torch.cuda.synchronize()
        for mode, (used, res) in results.items():
            same = torch.equal(res, A)
            if mode == "abi-split":   # the split sums owned and foreign columns separately: reduction-order tolerance
                err = float(((res - A).abs() / A.abs().clamp_min(1e-300)).max())
                print("mode %-10s (ran as %-6s) world %d grid %d^3: max relative difference to the 1-GPU result %.2e" % (mode, used, world, g, err))
                ok &= err < 1e-9
                continue
            print("mode %-10s (ran as %-6s) world %d grid %d^3: bitwise equal to 1-GPU result: %s" % (mode, used, world, g, same))
            ok &= same
        if g <= 160:
            import oracle
            x0 = torch.empty(rows, dtype=torch.float64, device="cuda")
            ctx.call("cme213_fill_uniform_f64", x0, rows, 9)
            want = oracle.csr_iterate(fp.cpu().numpy(), fc.cpu().numpy(), fv.cpu().numpy(), x0.cpu().numpy(), a.iters, omp=True)
            same = np.array_equal(A.cpu().numpy(), want)
            print("1-GPU result bitwise equal to the CPU oracle: %s" % same)
            ok &= same
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    dist.barrier()
    dist.destroy_process_group()
    if not int(flag.item()):
        sys.exit(1)


if __name__ == "__main__":
    main()
