"""CPU tests (gloo, world_size 2 and 3) of the row-sharded iterate: partition, halo/full exchange and
ping-pong logic of 2012-04_stanford_cme213_b200/sharded.py.  The per-rank SpMV is injected (the oracle's
CSR loop on numpy views), so no GPU is involved; on the GPU box the same class is bound to the CUDA
kernel (bench.py, tests/test_sharded_gpu.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import pkg


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, grid, mode, iters, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import importlib
        import oracle
        from oracle import generators as g
        sh = importlib.import_module("2012-04_stanford_cme213_b200.sharded")
        nx, ny, nz = grid
        rows = nx * ny * nz
        lo, hi = sh.row_partition(rows, world)[rank]
        ptr, col, val = g.laplace3d(nx, ny, nz, lo, hi)

        def spmv(a, b, x_full, y_full):  # rows [a,b) global; shard arrays are local
            if b <= a:
                return
            p = ptr[a - lo:b - lo + 1]
            y = oracle.csr_spmv(p - p[0], col[p[0]:p[-1]], val[p[0]:p[-1]], x_full.numpy())
            y_full.numpy()[a:b] = y

        x0 = np.random.default_rng(1).uniform(0, 1, rows)
        xa, xb = torch.from_numpy(x0.copy()), torch.zeros(rows, dtype=torch.float64)
        it = sh.ShardedIterate(rows, world, rank, spmv, xa, xb, halo=nx * ny, mode=mode, dist=dist)
        it.iterate(iters)
        full = it.gather_result(0).numpy()
        if rank == 0:
            fp, fc, fv = g.laplace3d(nx, ny, nz)
            want = oracle.csr_iterate(fp, fc, fv, x0, iters)
            out.put((it.mode, bool(np.array_equal(full, want)), it.comm_bytes_per_step))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,grid,mode,expect_mode", [
    (2, (6, 5, 8), "halo", "halo"),
    (2, (6, 5, 8), "full", "full"),
    (2, (4, 4, 3), "halo", "full"),    # rows not divisible by 2 planes each: ragged -> falls back... (48 rows / 2 ok, 1.5 planes)
    (3, (5, 4, 9), "halo", "halo"),
    (3, (5, 4, 6), "halo", "halo"),    # shards of exactly two planes: boundary rows cover the whole shard
    (2, (6, 5, 8), "p2p", "p2p"),
    (3, (5, 4, 9), "p2p", "p2p"),
])
def test_sharded_iterate_matches_single_process(world, grid, mode, expect_mode):
    ctx = mp.get_context("spawn")
    out = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, grid, mode, 4, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    got_mode, equal, comm = out.get()
    assert equal, "sharded result differs from the single-process oracle"
    rows = grid[0] * grid[1] * grid[2]
    assert got_mode == expect_mode
    if got_mode == "halo":
        assert comm == (world - 1) * 2 * grid[0] * grid[1] * 8
    elif got_mode == "p2p":
        assert comm == grid[0] * grid[1] * 8   # rank 0 has one neighbour
    else:
        assert comm == (world - 1) * (rows // world) * 8


def test_row_partition_and_bandwidth():
    sh = __import__("importlib").import_module("2012-04_stanford_cme213_b200.sharded")
    assert sh.row_partition(10, 3) == [(0, 4), (4, 8), (8, 10)]  # ceil(rows/P) each, last takes the rest (2dHeat.cpp:284-306)
    assert sh.row_partition(8, 4) == [(0, 2), (2, 4), (4, 6), (6, 8)]
    assert sh.row_partition(3, 5) == [(0, 1), (1, 2), (2, 3), (3, 3), (3, 3)]
    from oracle import generators as g
    ptr, col, _ = g.laplace3d(5, 4, 3)
    assert sh.half_bandwidth(torch.from_numpy(ptr), torch.from_numpy(col)) == 20
    p2, c2, _ = g.laplace3d(5, 4, 3, 20, 40)
    assert sh.half_bandwidth(torch.from_numpy(p2), torch.from_numpy(c2), 20) == 20


# ---- fused exchange on CPU: torch shared-memory tensors stand in for the CUDA-IPC peer mappings --------------------
class _FakePeers:
    """PeerVectors look-alike: prev(which) / next(which) give the neighbour's full-length vector (a tensor that every
    process maps) instead of a device address."""

    def __init__(self, vectors, rank, world):
        self.v, self.rank, self.world = vectors, rank, world

    def prev(self, which):
        return self.v[self.rank - 1][which] if self.rank > 0 else None

    def next(self, which):
        return self.v[self.rank + 1][which] if self.rank < self.world - 1 else None


def _fused_worker(rank, world, port, grid, iters, vectors, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import importlib
        import oracle
        from oracle import generators as g
        sh = importlib.import_module("2012-04_stanford_cme213_b200.sharded")
        nx, ny, nz = grid
        rows = nx * ny * nz
        lo, hi = sh.row_partition(rows, world)[rank]
        ptr, col, val = g.laplace3d(nx, ny, nz, lo, hi)

        def rows_of(a, b, x_full):
            p = ptr[a - lo:b - lo + 1]
            return oracle.csr_spmv(p - p[0], col[p[0]:p[-1]], val[p[0]:p[-1]], x_full.numpy())

        def spmv(a, b, x_full, y_full):
            if b > a:
                y_full.numpy()[a:b] = rows_of(a, b, x_full)

        def push(a, b, h, x_full, y_full, peer_prev, peer_next):
            # what cme213_spmv_csr_f64_halo_push does: both boundary planes, each stored locally and into the neighbour
            first, last = rows_of(a, a + h, x_full), rows_of(b - h, b, x_full)
            y_full.numpy()[a:a + h] = first
            y_full.numpy()[b - h:b] = last
            if peer_prev is not None:
                peer_prev.numpy()[a:a + h] = first
            if peer_next is not None:
                peer_next.numpy()[b - h:b] = last
        spmv.push = push

        xa, xb = vectors[rank]
        it = sh.ShardedIterate(rows, world, rank, spmv, xa, xb, halo=nx * ny, mode="fused", dist=dist,
                               peers=_FakePeers(vectors, rank, world))
        assert it.mode == "fused"
        it.iterate(iters)
        full = it.gather_result(0).numpy()
        if rank == 0:
            fp, fc, fv = g.laplace3d(nx, ny, nz)
            x0 = np.random.default_rng(1).uniform(0, 1, rows)
            out.put(bool(np.array_equal(full, oracle.csr_iterate(fp, fc, fv, x0, iters))))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,grid", [(2, (6, 5, 8)), (3, (5, 4, 9)), (3, (5, 4, 6))])
def test_fused_exchange_logic_matches_single_process(world, grid):
    rows = grid[0] * grid[1] * grid[2]
    x0 = np.random.default_rng(1).uniform(0, 1, rows)
    vectors = [(torch.from_numpy(x0.copy()).share_memory_(), torch.zeros(rows, dtype=torch.float64).share_memory_())
               for _ in range(world)]
    ctx = mp.get_context("spawn")
    out = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_fused_worker, args=(r, world, port, grid, 4, vectors, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert out.get(), "fused exchange result differs from the single-process oracle"


def test_partition_rows_host_function():
    """cme213_partition_rows is host code: bounds by nonzeros (binary search on the row pointers) or by rows."""
    import ctypes as C
    lib = pkg().load()
    b = (C.c_int64 * 5)()
    assert lib.cme213_partition_rows(1000, 4, None, 1, b) == 0 and list(b) == [0, 250, 500, 750, 1000]
    assert lib.cme213_partition_rows(10, 4, None, 1, b) == 0 and list(b) == [0, 3, 6, 9, 10]      # ceil(rows/P) each, the rest to the last
    lens = np.array([100] * 10 + [1] * 990)
    ptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    assert lib.cme213_partition_rows(1000, 2, C.c_void_p(ptr.ctypes.data), 1, b) == 0
    assert b[0] == 0 and b[2] == 1000 and abs(int(ptr[b[1]]) - int(ptr[-1]) // 2) <= 100
    assert lib.cme213_partition_rows(1000, 3, C.c_void_p(ptr.ctypes.data), 128, b) == 0
    assert all(v % 128 == 0 or v == 1000 for v in list(b)[:4]) and list(b)[:4] == sorted(list(b)[:4])
    assert lib.cme213_partition_rows(-1, 2, None, 1, b) != 0
