"""Row-sharded iterated y = A x across the GPUs of one box (config C5).

One process per GPU (torchrun); `torch.distributed` (NCCL over NVLink 5 / NVSwitch) carries the x
exchange, the SpMV itself is the library's CSR kernel on each rank's row block.

Decomposition: contiguous row blocks, ceil(rows/P) rows per rank with the last rank taking the
remainder -- the 1-D stripe decomposition of the reference's only distributed code
(hw/hw5/programming/2dHeat.cpp:284-306), with its start-transfers / compute-interior / wait /
compute-borders overlap (2dHeat.cpp:696-723) turned around for SpMV: the *boundary* rows are
computed first because they are what the neighbours need, the exchange then runs on a second
stream while the interior rows are computed.

Exchange modes
  full : every rank all-gathers its whole y slice into everybody's x (N*8/P bytes sent per rank).
         Always correct; costs more than the SpMV itself beyond 2 GPUs for the 7-point stencil.
  p2p  : same data as `halo`, moved by NCCL send/recv between neighbours straight out of / into the full-length
         vector (no pack / unpack copies).
  halo : for a matrix of half-bandwidth `halo` (max |col - row|), rank r only needs the last `halo`
         entries of rank r-1 and the first `halo` of rank r+1.  Each rank contributes its first and
         last `halo` entries to ONE all-gather (2*halo values per rank) and copies the two pieces it
         needs into its full-length x.  For the 512^3 grid that is 2 x 2 MiB per rank instead of
         1 GiB/P.

Row sharding does not change the order in which a row is summed, so every rank's slice is bitwise
identical to the single-GPU result.

The iteration skeleton is the ping-pong loop of hw/hw1/programming/pagerank.cu:59-67, 128-132.
"""
import math


def row_partition(rows, world):
    """[begin, end) of every rank: ceil(rows/world) rows each, the last ranks take what is left."""
    per = -(-rows // world)
    return [(min(r * per, rows), min((r + 1) * per, rows)) for r in range(world)]


def equal_partition_ok(rows, world):
    return rows % world == 0


class ShardedIterate:
    """Iterated y = A x for one rank's row block.

    spmv(row_lo, row_hi, x_full, y_full): computes rows [row_lo,row_hi) (GLOBAL indices inside this
    rank's block) of y_full from x_full.  On GPUs it is bound to cme213_spmv_csr_* by
    `cuda_spmv`; the CPU tests inject a numpy implementation to exercise the partition / exchange
    logic over gloo.
    """

    def __init__(self, rows, world, rank, spmv, x_a, x_b, halo=None, mode="halo", dist=None, streams=None):
        self.rows, self.world, self.rank = rows, world, rank
        self.spmv = spmv
        self.x = [x_a, x_b]  # full-length ping-pong vectors
        self.dist = dist
        self.parts = row_partition(rows, world)
        self.lo, self.hi = self.parts[rank]
        self.mode = mode if world > 1 else "none"
        self.halo = int(halo) if halo is not None else None
        if self.mode in ("halo", "p2p"):
            smallest = min(e - b for b, e in self.parts)
            if self.halo is None or not equal_partition_ok(rows, world) or smallest < 2 * self.halo:
                self.mode = "full"  # shards too thin (or ragged) for the two-piece halo layout
        if self.mode == "full" and not equal_partition_ok(rows, world):
            raise ValueError("full all-gather needs rows divisible by the world size (got %d / %d)" % (rows, world))
        self.compute_stream, self.comm_stream = streams if streams else (None, None)
        if self.mode == "halo":
            h = self.halo
            self.send = x_a.new_empty(2 * h)
            self.recv = x_a.new_empty(2 * h * world)
        self.comm_bytes_per_step = 0
        if self.mode == "full":
            self.comm_bytes_per_step = (world - 1) * (self.hi - self.lo) * x_a.element_size()
        elif self.mode == "halo":
            self.comm_bytes_per_step = (world - 1) * 2 * self.halo * x_a.element_size()
        elif self.mode == "p2p":
            self.comm_bytes_per_step = ((rank > 0) + (rank < world - 1)) * self.halo * x_a.element_size()

    # -- helpers ---------------------------------------------------------------------------------
    def _on(self, stream):
        import contextlib
        import torch
        return torch.cuda.stream(stream) if stream is not None else contextlib.nullcontext()

    def _event(self, stream):
        import torch
        if stream is None:
            return None
        e = torch.cuda.Event()
        e.record(stream)
        return e

    def step(self, src, dst):
        """x[dst] = A x[src] (own rows computed locally, the rest obtained by the exchange)."""
        xin, xout = self.x[src], self.x[dst]
        lo, hi = self.lo, self.hi
        if self.mode == "none":
            self.spmv(lo, hi, xin, xout)
            return
        if self.mode == "full":
            self.spmv(lo, hi, xin, xout)
            # in place: every rank's slice already sits at its final position in xout
            self.dist.all_gather_into_tensor(xout, xout[lo:hi])
            return
        h = self.halo
        cs, ms = self.compute_stream, self.comm_stream
        if self.mode == "p2p":
            with self._on(cs):
                # 1. boundary rows first (one launch when the kernel binding offers it): they are the neighbours' halo
                if hasattr(self.spmv, "two_ranges"):
                    self.spmv.two_ranges(lo, hi, h, xin, xout)
                else:
                    self.spmv(lo, lo + h, xin, xout)
                    self.spmv(hi - h, hi, xin, xout)
                ready = self._event(cs)
            with self._on(ms):
                # 2. neighbour exchange straight from / into the full-length vector: no pack, no unpack
                if ready is not None:
                    ms.wait_event(ready)
                D = self.dist
                ops = []
                if self.rank > 0:
                    ops += [D.P2POp(D.isend, xout[lo:lo + h], self.rank - 1), D.P2POp(D.irecv, xout[lo - h:lo], self.rank - 1)]
                if self.rank < self.world - 1:
                    ops += [D.P2POp(D.isend, xout[hi - h:hi], self.rank + 1), D.P2POp(D.irecv, xout[hi:hi + h], self.rank + 1)]
                for w in D.batch_isend_irecv(ops):
                    w.wait()
                done = self._event(ms)
            with self._on(cs):
                self.spmv(lo + h, hi - h, xin, xout)
                if done is not None:
                    cs.wait_event(done)
            return
        with self._on(cs):
            # 1. boundary rows first: they are the neighbours' halo
            if hasattr(self.spmv, "two_ranges"):
                self.spmv.two_ranges(lo, hi, h, xin, xout)
            else:
                self.spmv(lo, lo + h, xin, xout)
                self.spmv(hi - h, hi, xin, xout)
            self.send[:h].copy_(xout[lo:lo + h])
            self.send[h:].copy_(xout[hi - h:hi])
            ready = self._event(cs)
        with self._on(ms):
            # 2. one all-gather of the boundary pieces, overlapped with the interior rows
            if ready is not None:
                ms.wait_event(ready)
            self.dist.all_gather_into_tensor(self.recv, self.send)
            if self.rank > 0:
                xout[lo - h:lo].copy_(self.recv[(self.rank - 1) * 2 * h + h:(self.rank - 1) * 2 * h + 2 * h])
            if self.rank < self.world - 1:
                xout[hi:hi + h].copy_(self.recv[(self.rank + 1) * 2 * h:(self.rank + 1) * 2 * h + h])
            done = self._event(ms)
        with self._on(cs):
            # 3. interior rows need no remote data
            self.spmv(lo + h, hi - h, xin, xout)
            if done is not None:
                cs.wait_event(done)

    def capture_pair(self):
        """Records step(0,1) + step(1,0) -- kernels, copies and the NCCL all-gather -- into one CUDA graph, so that
        a pair of passes costs one launch from the host.  With 8 GPUs a pass takes ~0.35 ms and the ~10 host-side
        launches per pass would otherwise be the bottleneck.  Returns a callable that replays the pair, or None
        when capture is not possible (CPU tensors, capture error): callers then keep calling step()."""
        import torch
        if self.compute_stream is None and self.world > 1:
            return None
        if not self.x[0].is_cuda:
            return None
        try:
            # one eager pair first: creates communicators, CTA-order tables and workspaces outside the capture
            self.step(0, 1)
            self.step(1, 0)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            side = torch.cuda.Stream()
            saved = (self.compute_stream, self.comm_stream)
            with torch.cuda.graph(graph, stream=side):
                # inside the capture the "compute" stream is the capturing stream
                if self.world > 1:
                    self.compute_stream = torch.cuda.current_stream()
                self.step(0, 1)
                self.step(1, 0)
            self.compute_stream, self.comm_stream = saved
            torch.cuda.synchronize()
            return graph.replay
        except Exception as e:  # noqa: BLE001 -- any capture problem means "run eagerly"
            import sys
            print("[sharded] CUDA graph capture unavailable, running eagerly: %r" % (e,), file=sys.stderr)
            try:
                self.compute_stream, self.comm_stream = saved
            except Exception:
                pass
            torch.cuda.synchronize()
            return None

    def iterate(self, nr_iterations):
        """nr_iterations (even, as pagerank.cu:61 asserts) passes; the result ends in x[0]."""
        if nr_iterations % 2 != 0:
            raise ValueError("nr_iterations must be even")
        for _ in range(0, nr_iterations, 2):
            self.step(0, 1)
            self.step(1, 0)

    def gather_result(self, which=0):
        """Every rank's slice of x[which] in one full vector (for checks; not part of the timed path)."""
        x = self.x[which]
        if self.world == 1:
            return x
        if equal_partition_ok(self.rows, self.world):
            out = x.new_empty(self.rows)
            self.dist.all_gather_into_tensor(out, x[self.lo:self.hi].contiguous())
            return out
        pieces = [x.new_empty(e - b) for b, e in self.parts]
        self.dist.all_gather(pieces, x[self.lo:self.hi].contiguous())
        import torch
        return torch.cat(pieces)


def grid3d_block_order(nx, ny, nz, row_lo, row_hi, block_rows=256, strip_lines=8, device="cuda"):
    """CTA order for cme213_spmv_csr_*_ordered on rows [row_lo,row_hi) of a natural-order nx*ny*nz grid operator:
    row blocks sorted by (strip of `strip_lines` grid lines in y, z, position inside the strip), so consecutive
    CTAs walk z fastest inside a thin strip and the three x planes a 7-point row touches stay in L2."""
    import torch
    nblk = -(-(row_hi - row_lo) // block_rows)
    first = torch.arange(nblk, device=device, dtype=torch.int64) * block_rows + row_lo
    y = (first // nx) % ny
    z = first // (nx * ny)
    key = ((y // strip_lines) * nz + z) * (nblk + 1) + torch.arange(nblk, device=device, dtype=torch.int64) % (nblk + 1)
    return torch.argsort(key, stable=True).to(torch.int32)


def cuda_spmv(ctx, row_lo_global, d_row_ptr_local, d_col, d_val, kernel=0, order_fn=None):
    """Binds the C-ABI CSR kernel to a rank's shard.  The shard's row_ptr is rebased to 0 and indexed by
    local row, its column indices are global; y/x are the full-length vectors, so the kernel is handed
    y shifted by the shard's first row."""
    import torch
    fn = "cme213_spmv_csr_f64" if d_val.dtype == torch.float64 else "cme213_spmv_csr_f32"
    esize = d_val.element_size()

    orders = {}

    def spmv(lo, hi, x_full, y_full):
        if hi <= lo:
            return
        if order_fn is not None:
            # one CTA order per launch range (whole shard, boundary pieces, interior), built on first use
            if (lo, hi) not in orders:
                orders[(lo, hi)] = order_fn(lo, hi)
            ctx.call(fn + "_ordered", lo - row_lo_global, hi - row_lo_global, d_row_ptr_local, d_col, d_val, x_full,
                     y_full.data_ptr() + row_lo_global * esize, orders[(lo, hi)], orders[(lo, hi)].numel(),
                     stream=torch.cuda.current_stream())
            return
        ctx.call(fn, lo - row_lo_global, hi - row_lo_global, d_row_ptr_local, d_col, d_val, x_full,
                 y_full.data_ptr() + row_lo_global * esize, kernel, stream=torch.cuda.current_stream())

    def spmv_two_ranges(lo, hi, h, x_full, y_full):
        """rows [lo, lo+h) and [hi-h, hi) of the shard in ONE launch (h a multiple of the kernel's block rows)."""
        key = ("pair", lo, hi, h)
        if key not in orders:
            br = int(ctx.lib.cme213_spmv_csr_block_rows())
            if h % br or (hi - lo) % br:
                orders[key] = None
            else:
                nb, hb = (hi - lo) // br, h // br
                orders[key] = torch.cat([torch.arange(hb, device="cuda"), torch.arange(nb - hb, nb, device="cuda")]).to(torch.int32)
        o = orders[key]
        if o is None:
            spmv(lo, lo + h, x_full, y_full)
            spmv(hi - h, hi, x_full, y_full)
            return
        ctx.call(fn + "_ordered", lo - row_lo_global, hi - row_lo_global, d_row_ptr_local, d_col, d_val, x_full,
                 y_full.data_ptr() + row_lo_global * esize, o, o.numel(), stream=torch.cuda.current_stream())

    spmv.two_ranges = spmv_two_ranges
    return spmv


def half_bandwidth(row_ptr, col, row_offset=0):
    """max |col - row| of a CSR block (torch tensors, any device); rows are row_offset + local index."""
    import torch
    n = row_ptr.numel() - 1
    best = 0
    step = 1 << 22
    for b in range(0, n, step):
        e = min(b + step, n)
        counts = (row_ptr[b + 1:e + 1] - row_ptr[b:e]).to(torch.int64)
        rows = torch.repeat_interleave(torch.arange(b, e, device=row_ptr.device) + row_offset, counts)
        c = col[int(row_ptr[b]):int(row_ptr[e])].to(torch.int64)
        if c.numel():
            best = max(best, int((c - rows).abs().max()))
    return best
