"""numpy restatement of the device-side input generators (csrc/generate.cu, csrc/common.cuh hash32)
plus small host-side builders used by the tests.  TEST INFRASTRUCTURE ONLY."""
import numpy as np

M32 = np.uint64(0xFFFFFFFF)


def hash32(seed, idx):
    """Bit-for-bit numpy version of cme213::hash32 (common.cuh)."""
    idx = np.asarray(idx, dtype=np.uint64)
    h = np.full(idx.shape, (int(seed) ^ 0x9E3779B9) & 0xFFFFFFFF, dtype=np.uint64)
    h ^= idx & M32
    h = (h * np.uint64(0x85EBCA6B)) & M32
    h ^= h >> np.uint64(13)
    h ^= ((idx >> np.uint64(32)) * np.uint64(0xC2B2AE35)) & M32
    h = (h * np.uint64(0xC2B2AE35)) & M32
    h ^= h >> np.uint64(16)
    h = (h * np.uint64(0x27D4EB2F)) & M32
    h ^= h >> np.uint64(15)
    return h.astype(np.uint32)


def uniform_pm1_f32(h):
    v = (h >> np.uint32(8)).astype(np.float32) * np.float32(1.1920928955078125e-07) - np.float32(1.0)
    v[v == 0] = np.float32(1.1920928955078125e-07)
    return v.astype(np.float32)


def uniform_pm1_f64(h):
    v = h.astype(np.float64) * 4.656612873077392578125e-10 - 1.0
    v[v == 0] = 4.656612873077392578125e-10
    return v


def fill_uniform_f32(n, seed):
    return uniform_pm1_f32(hash32(seed, np.arange(n, dtype=np.uint64)))


def fill_uniform_f64(n, seed):
    return uniform_pm1_f64(hash32(seed, np.arange(n, dtype=np.uint64)))


def fill_hash_u32(n, seed, modulus=0):
    h = hash32(seed, np.arange(n, dtype=np.uint64))
    return (h % np.uint32(modulus)).astype(np.uint32) if modulus else h


def random_flags(n, seed, prob_num):
    """Head flags as a bool array (bit 0 always set), as cme213_fill_random_flags produces them."""
    f = hash32(seed, np.arange(n, dtype=np.uint64)) < np.uint32(prob_num)
    if n:
        f[0] = True
    return f


def pack_flags(flags):
    """bool[n] -> uint32 words, bit i of word i/32."""
    n = len(flags)
    words = (n + 31) // 32
    padded = np.zeros(words * 32, dtype=np.uint8)
    padded[:n] = flags
    return np.packbits(padded.reshape(-1, 32)[:, ::-1], axis=1).view(">u4").astype(np.uint32).ravel()


def offsets_to_flags(s, n):
    f = np.zeros(n, dtype=bool)
    s = np.asarray(s)
    f[s[s < n]] = True
    return f


def flags_to_offsets(flags):
    """Segment start offsets in the reference's layout: s[0]=0, ..., s[p-1]=n."""
    n = len(flags)
    return np.concatenate([np.nonzero(flags)[0], [n]]).astype(np.int32)


def laplace2d(nx, ny, dtype=np.float32):
    """5-point Laplacian, natural order, sorted columns, diag 4 / off -1 (cme213_gen_laplace2d_f32)."""
    ptr, col, val = [0], [], []
    for iy in range(ny):
        for ix in range(nx):
            i = iy * nx + ix
            if iy > 0: col.append(i - nx); val.append(-1)
            if ix > 0: col.append(i - 1); val.append(-1)
            col.append(i); val.append(4)
            if ix < nx - 1: col.append(i + 1); val.append(-1)
            if iy < ny - 1: col.append(i + nx); val.append(-1)
            ptr.append(len(col))
    return np.array(ptr, np.int32), np.array(col, np.int32), np.array(val, dtype)


def laplace3d(nx, ny, nz, row_begin=0, row_end=None, dtype=np.float64):
    """7-point Laplacian rows [row_begin,row_end), global columns, rebased row pointers."""
    n = nx * ny * nz
    row_end = n if row_end is None else row_end
    i = np.arange(row_begin, row_end, dtype=np.int64)
    ix, iy, iz = i % nx, (i // nx) % ny, i // (nx * ny)
    offs = [(-nx * ny, iz > 0), (-nx, iy > 0), (-1, ix > 0), (0, np.ones_like(i, bool)), (1, ix < nx - 1),
            (nx, iy < ny - 1), (nx * ny, iz < nz - 1)]
    cols = np.stack([i + o for o, _ in offs], axis=1)
    mask = np.stack([m for _, m in offs], axis=1)
    vals = np.where(np.array([o for o, _ in offs]) == 0, 6.0, -1.0)[None, :].repeat(len(i), 0)
    ptr = np.concatenate([[0], np.cumsum(mask.sum(1))]).astype(np.int32)
    return ptr, cols[mask].astype(np.int32), vals[mask].astype(dtype)


def powerlaw_lengths(rows, r0, num, den, cap, seed):
    h = hash32(seed, np.arange(rows, dtype=np.uint64)).astype(np.uint64)
    delta = (np.uint64(num) * (M32 - h)) // (np.uint64(den) * (h + np.uint64(1)))
    t = np.minimum(np.uint64(r0) + delta, np.uint64(cap))
    return np.concatenate([[0], np.cumsum(t)]).astype(np.int32)


def random_csr_fill(ptr, num_cols, window, seed):
    ptr = np.asarray(ptr, dtype=np.int64)
    rows, nnz = len(ptr) - 1, int(ptr[-1])
    j = np.arange(nnz, dtype=np.int64)
    r = np.searchsorted(ptr, j, side="right") - 1
    length = (ptr[1:] - ptr[:-1])[r]
    pos = j - ptr[r]
    W = window if 0 < window < num_cols else num_cols
    base = np.clip(r - W // 2, 0, num_cols - W)
    stride = (W // length).astype(np.uint64)
    h = hash32(seed, j.astype(np.uint64)).astype(np.uint64)
    c = base.astype(np.uint64) + pos.astype(np.uint64) * stride + ((h * stride) >> np.uint64(32))
    c = np.minimum(c, np.uint64(num_cols - 1))
    val = uniform_pm1_f32(hash32(seed ^ 0x5bd1e995, j.astype(np.uint64)))
    return c.astype(np.int32), val


def pagerank_graph(n, avg_edges, seed):
    """pagerank.cu:185-204 with hash32 in place of rand()."""
    i = np.arange(n, dtype=np.int64)
    length = (i % (2 * avg_edges - 1) + 1)
    indices = np.concatenate([[0], np.cumsum(length)]).astype(np.uint32)
    nnz = int(indices[-1])
    edges = (hash32(seed, np.arange(nnz, dtype=np.uint64)) % np.uint32(n)).astype(np.uint32)
    inv = (np.float32(1.0) / length.astype(np.float32)).astype(np.float32)
    nodes = np.full(n, np.float32(1.0) / np.float32(n), dtype=np.float32)
    return indices, edges, inv, nodes


def random_csr(rows, cols, mean_len, seed, dtype=np.float32, max_len=None):
    """Small general CSR matrix for tests: Poisson-ish row lengths (some empty), random sorted columns."""
    rng = np.random.default_rng(seed)
    length = rng.poisson(mean_len, rows)
    if max_len is not None:
        length = np.minimum(length, max_len)
    length = np.minimum(length, cols)
    ptr = np.concatenate([[0], np.cumsum(length)]).astype(np.int32)
    col = np.concatenate([np.sort(rng.choice(cols, l, replace=False)) for l in length] + [np.zeros(0, np.int64)]).astype(np.int32)
    val = rng.uniform(-1, 1, int(ptr[-1])).astype(dtype)
    val[val == 0] = 0.5
    return ptr, col, val


def fp_inputs(n, p, q, seed, dtype=np.float32):
    """(a, s, k, x) with the semantics of readMM.py:31-60: a,x ~ U(-1,1), sorted random cut points, k ~ U[0,q)."""
    rng = np.random.default_rng(seed)
    a = rng.uniform(-1, 1, n).astype(dtype)
    x = rng.uniform(-1, 1, q).astype(dtype)
    cuts = np.sort(rng.choice(np.arange(1, n), size=max(p - 2, 0), replace=False)) if n > 1 else np.zeros(0, np.int64)
    s = np.concatenate([[0], cuts, [n]]).astype(np.int32)
    k = rng.integers(0, q, n).astype(np.int32)
    return a, s, k, x



def segment_offsets(n, num_cuts, seed):
    """cme213_gen_segment_offsets restated: readMM.py:44-49's sorted random cut points, duplicates dropped."""
    cuts = np.uint64(1) + hash32(seed, np.arange(num_cuts, dtype=np.uint64)).astype(np.uint64) % np.uint64(max(n - 1, 1))
    return np.concatenate([[0], np.unique(cuts), [n]]).astype(np.int32)


def coo_to_csr(num_rows, row, col, val):
    """cme213_coo_to_csr_* restated: stable sort by (row, column), duplicates kept in input order; what scipy's
    coo -> csr does for the MatrixMarket inputs of readMM.py:18-24 (without summing duplicates)."""
    row, col = np.asarray(row, dtype=np.int64), np.asarray(col, dtype=np.int64)
    order = np.lexsort((col, row))          # last key is the primary one; lexsort is stable
    ptr = np.zeros(num_rows + 1, dtype=np.int64)
    np.add.at(ptr, row + 1, 1)
    return np.cumsum(ptr).astype(np.int32), col[order].astype(np.int32), np.asarray(val)[order]
