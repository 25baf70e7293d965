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
  fused: the boundary launch itself stores the halo planes into the neighbours' vectors (peer memory over NVLink,
         mapped with CUDA IPC: cme213_spmv_csr_f64_halo_push); NCCL only orders the ranks afterwards (a one-element
         all-reduce on the second stream, hidden behind the interior rows).  Needs vectors from `shared_vector`.
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


def partition_rows(lib, rows, world, h_row_ptr=None, align=128):
    """Row-block bounds (world + 1 entries) from the C ABI: balanced by nonzeros when the host row pointers (numpy
    int32) are given, ceil(rows/world) rows each otherwise; rounded to `align` rows."""
    import ctypes as C
    out = (C.c_int64 * (world + 1))()
    p = None if h_row_ptr is None else C.c_void_p(h_row_ptr.ctypes.data)
    if lib.cme213_partition_rows(rows, world, p, align, out):
        raise RuntimeError(lib.cme213_last_error().decode())
    return list(out)


class Shard:
    """One rank's cme213_shard (csrc/sharded.cu): the row-sharded iterate behind the C ABI.  This class only moves the
    export handles between the ranks (torch.distributed, any backend) and wraps the two device vectors; partition,
    exchange protocol, kernels and ordering all live in the library -- a C++ host does the same with MPI or a pipe."""

    def __init__(self, ctx, dist, rank, world, rows, bounds, d_row_ptr, d_col, d_val, mode=0):
        import ctypes as C
        import torch
        self.ctx, self.lib, self.rank, self.world, self.rows = ctx, ctx.lib, rank, world, rows
        self.keep = (d_row_ptr, d_col, d_val)
        b = (C.c_int64 * (world + 1))(*bounds)
        h = C.c_void_p()
        self._check(self.lib.cme213_shard_create_f64(ctx.h, rank, world, rows, b, C.c_void_p(d_row_ptr.data_ptr()),
                                                     C.c_void_p(d_col.data_ptr()), C.c_void_p(d_val.data_ptr()), mode, C.byref(h)))
        self.h = h
        self.lo, self.hi = bounds[rank], bounds[rank + 1]
        if world > 1:
            n = self.lib.cme213_shard_handle_bytes()
            blob = C.create_string_buffer(n)
            rc = self.lib.cme213_shard_export(self.h, blob)
            mine = None if rc else blob.raw
            err = self.lib.cme213_last_error().decode() if rc else None
            everyone = [None] * world
            dist.all_gather_object(everyone, mine)          # always reached: no rank is left waiting
            if any(e is None for e in everyone):
                raise RuntimeError(err or "a peer could not export its shard")
            rc = self.lib.cme213_shard_connect(self.h, C.create_string_buffer(b"".join(everyone), n * world))
            err = self.lib.cme213_last_error().decode() if rc else None
            oks = [None] * world
            dist.all_gather_object(oks, 0 if rc else 1)     # every rank learns whether all of them connected
            if not all(oks):
                raise RuntimeError(err or "a peer could not connect")
        self.mode = int(self.lib.cme213_shard_mode(self.h))
        self.comm_bytes_per_step = int(self.lib.cme213_shard_comm_bytes_per_step(self.h))
        self.x = [torch.as_tensor(_RawCuda(self.lib.cme213_shard_vector(self.h, w), rows, "<f8"), device="cuda") for w in (0, 1)]

    def _check(self, rc):
        if rc:
            raise RuntimeError("cme213 error %d: %s" % (rc, self.lib.cme213_last_error().decode()))

    def _stream(self, stream):
        import ctypes as C
        import torch
        st = torch.cuda.current_stream() if stream is None else stream
        return C.c_void_p(st.cuda_stream)

    def step(self, src, stream=None):
        self._check(self.lib.cme213_shard_step(self.h, src, self._stream(stream)))

    def iterate(self, nr_iterations, stream=None):
        self._check(self.lib.cme213_shard_iterate(self.h, nr_iterations, self._stream(stream)))

    def finish(self, stream=None):
        self._check(self.lib.cme213_shard_finish(self.h, self._stream(stream)))

    def close(self):
        if getattr(self, "h", None):
            self.x = None
            self.lib.cme213_shard_destroy(self.h)
            self.h = None


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

    def __init__(self, rows, world, rank, spmv, x_a, x_b, halo=None, mode="halo", dist=None, streams=None, peers=None):
        self.rows, self.world, self.rank = rows, world, rank
        self.spmv = spmv
        self.x = [x_a, x_b]  # full-length ping-pong vectors
        self.dist = dist
        self.parts = row_partition(rows, world)
        self.lo, self.hi = self.parts[rank]
        self.mode = mode if world > 1 else "none"
        self.halo = int(halo) if halo is not None else None
        self.peers = peers  # fused mode: PeerVectors of x_a / x_b
        if self.mode == "fused" and (peers is None or not hasattr(spmv, "push")):
            raise ValueError("fused exchange needs peer-mapped vectors (shared_vector / PeerVectors) and a CUDA spmv binding")
        if self.mode in ("halo", "p2p", "fused"):
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
        elif self.mode in ("p2p", "fused"):
            self.comm_bytes_per_step = ((rank > 0) + (rank < world - 1)) * self.halo * x_a.element_size()
        if self.mode == "fused":
            self.token = x_a.new_zeros(1)

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
        if self.mode == "fused":
            with self._on(cs):
                # 1. boundary rows, pushed into the neighbours' xout by the kernel's own stores
                self.spmv.push(lo, hi, h, xin, xout, self.peers.prev(dst), self.peers.next(dst))
                ready = self._event(cs)
            with self._on(ms):
                # 2. order the ranks: after this collective every neighbour's boundary launch has completed, so what
                #    it pushed into this rank's xout is there
                if ready is not None:
                    ms.wait_event(ready)
                self.dist.all_reduce(self.token)
                done = self._event(ms)
            with self._on(cs):
                self.spmv(lo + h, hi - h, xin, xout)
                if done is not None:
                    cs.wait_event(done)
            return
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

    def spmv_push(lo, hi, h, x_full, y_full, peer_prev, peer_next):
        """boundary launch with the halo exchange fused in; peer_prev / peer_next are the base addresses of the
        neighbours' output vectors (0 = no neighbour on that side)."""
        key = ("pair", lo, hi, h)
        if key not in orders:
            br = int(ctx.lib.cme213_spmv_csr_block_rows())
            if h % br or (hi - lo) % br:
                raise ValueError("fused exchange needs halo and shard sizes that are multiples of %d rows" % br)
            nb, hb = (hi - lo) // br, h // br
            orders[key] = torch.cat([torch.arange(hb, device="cuda"), torch.arange(nb - hb, nb, device="cuda")]).to(torch.int32)
        o = orders[key]
        if o is None:
            raise ValueError("fused exchange needs halo and shard sizes that are multiples of the kernel's block rows")
        shift = row_lo_global * esize
        ctx.call("cme213_spmv_csr_f64_halo_push", lo - row_lo_global, hi - row_lo_global, d_row_ptr_local, d_col, d_val, x_full,
                 y_full.data_ptr() + shift, o, o.numel(), h, (peer_prev + shift) if peer_prev else None,
                 (peer_next + shift) if peer_next else None, stream=torch.cuda.current_stream())

    spmv.two_ranges = spmv_two_ranges
    if d_val.dtype == torch.float64:
        spmv.push = spmv_push
    return spmv


class _RawCuda:
    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 3}


def shared_vector(lib, n, dtype="f8"):
    """A device vector from one cudaMalloc of its own (exportable to the other ranks), as (torch tensor, address)."""
    import ctypes as C
    import torch
    p = C.c_void_p()
    rc = lib.cme213_device_malloc(n * int(dtype[1:]), C.byref(p))
    if rc:
        raise RuntimeError(lib.cme213_last_error().decode())
    return torch.as_tensor(_RawCuda(p.value, n, "<" + dtype), device="cuda"), p.value


class PeerVectors:
    """The two ping-pong vectors of the neighbouring ranks, mapped into this process through CUDA IPC."""

    def __init__(self, lib, dist, rank, world, addrs):
        import ctypes as C
        nbytes = lib.cme213_ipc_handle_bytes()
        mine, failed = [], None
        for a in addrs:
            buf = C.create_string_buffer(nbytes)
            if lib.cme213_ipc_export(C.c_void_p(a), buf):
                failed = lib.cme213_last_error().decode()
            mine.append(buf.raw)
        everyone = [None] * world
        dist.all_gather_object(everyone, None if failed else mine)   # always reached: no rank is left waiting
        if failed or any(e is None for e in everyone):
            raise RuntimeError(failed or "a peer could not export its vectors")
        self.lib = lib
        self._prev, self._next = [0, 0], [0, 0]

        def open_(raw):
            p = C.c_void_p()
            if lib.cme213_ipc_open(C.create_string_buffer(raw, nbytes), C.byref(p)):
                raise RuntimeError(lib.cme213_last_error().decode())
            return p.value

        if rank > 0:
            self._prev = [open_(h) for h in everyone[rank - 1]]
        if rank < world - 1:
            self._next = [open_(h) for h in everyone[rank + 1]]

    def prev(self, which):
        return self._prev[which]

    def next(self, which):
        return self._next[which]


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
