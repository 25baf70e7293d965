"""Builds libcme213_b200.so in-tree with nvcc for sm_100a (no JIT cache, no torch extension).

Usage: python build.py [--force] [--verbose] [--tuning]   (--tuning compiles the environment-variable tuning hooks in)
The shared object sits next to this file so that it travels with the repository snapshot.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libcme213_b200.so")
SOURCES = ["ctx.cu", "transpose.cu", "scan.cu", "spmv_csr.cu", "spmv_ell.cu", "spmv_dia.cu", "convert.cu", "radix_sort.cu", "coo_to_csr.cu", "generate.cu", "sharded.cu",
           "host_api.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC,-fvisibility=hidden,-Wall", "-Xptxas", "-v"]


def source_hash():
    """sha1 over csrc/, host-independent: the library carries it (cme213_build_id) and capi.load() compares, so a
    stale git-ignored .so cannot be used silently against newer sources."""
    import hashlib
    h = hashlib.sha1()
    for root in (CSRC, os.path.join(HERE, "..", "include")):
        for f in sorted(os.listdir(root)):
            if f.endswith((".cu", ".cuh", ".h")):
                h.update(f.encode())
                with open(os.path.join(root, f), "rb") as fh:
                    h.update(fh.read())
    return h.hexdigest()[:16]


def _compile(src, verbose, extra=()):
    obj = os.path.join(OBJ, src.replace(".cu", ".o"))
    cmd = ["nvcc"] + NVCC_FLAGS + list(extra) + ["-c", os.path.join(CSRC, src), "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed on " + src)
    with open(obj + ".ptxas.txt", "w") as f:
        f.write(r.stderr)
    return obj


def built_hash():
    try:
        with open(os.path.join(OBJ, "source_hash.txt")) as f:
            return f.read().strip()
    except OSError:
        return None


def build(force=False, verbose=False, tuning=False):
    sources = [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    want = source_hash() + ("-tuning" if tuning else "")
    if not force and os.path.exists(LIB) and built_hash() == want:
        return LIB
    if subprocess.run(["which", "nvcc"], capture_output=True).returncode != 0:
        raise RuntimeError("libcme213_b200.so is missing or older than csrc/ (built %s, sources %s) and nvcc is not here to "
                           "rebuild it" % (built_hash(), want))
    os.makedirs(OBJ, exist_ok=True)
    extra = ['-DCME213_SOURCE_HASH="%s"' % want] + (["-DCME213_TUNING"] if tuning else [])
    with ThreadPoolExecutor(max_workers=min(8, len(sources))) as ex:
        objs = list(ex.map(lambda s: _compile(s, verbose, extra), sources))
    cmd = ["nvcc", "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("link failed")
    with open(os.path.join(OBJ, "source_hash.txt"), "w") as f:
        f.write(want + "\n")
    return LIB


DRIVERS = {"fp": ["fp_main.cpp", "matrix_io.cpp"], "pagerank": ["pagerank_main.cpp"],
           "check_output": ["check_output_main.cpp", "matrix_io.cpp"], "radixsort": ["radixsort_main.cpp"],
           "csr_iterate": ["csr_iterate_main.cpp"]}


def build_drivers(force=False):
    """The C++ drop-in drivers (host/), linked against the in-tree shared library."""
    host, bindir = os.path.join(HERE, "host"), os.path.join(HERE, "bin")
    os.makedirs(bindir, exist_ok=True)
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    for name, srcs in DRIVERS.items():
        out = os.path.join(bindir, name)
        paths = [os.path.join(host, s) for s in srcs]
        newest = max(os.path.getmtime(p) for p in paths + [os.path.join(host, "matrix_io.h"), LIB])
        if not force and os.path.exists(out) and os.path.getmtime(out) >= newest:
            continue
        cmd = [gxx, "-O2", "-fopenmp", "-std=c++17", "-o", out] + paths + ["-L" + HERE, "-lcme213_b200", "-Wl,-rpath,$ORIGIN/.."]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
            raise RuntimeError("driver build failed: " + name)


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv, tuning="--tuning" in sys.argv))
    build_drivers(force="--force" in sys.argv)
