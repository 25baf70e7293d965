"""CPU tests of the drop-in boundary: the shared library builds/loads, exports every symbol that
include/cme213_b200.h declares, the ctypes table covers them all, and -- without a GPU -- the
product path fails loudly instead of falling back to anything."""
import os
import re
import subprocess

import pytest

from conftest import ROOT, pkg


def header_symbols():
    text = open(os.path.join(ROOT, "include", "cme213_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"\b(cme213_[a-z0-9_]+|device_graph_iterate)\s*\(", text)
    return sorted(set(names))


def test_header_declares_expected_surface():
    syms = header_symbols()
    assert "device_graph_iterate" in syms and "cme213_spmv_csr_f64" in syms and len(syms) >= 45


def test_library_exports_every_declared_symbol():
    p = pkg()
    lib = p.load()
    out = subprocess.run(["nm", "-D", "--defined-only", p.capi.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(line.split()[-1] for line in out.splitlines() if line.strip())
    for s in header_symbols():
        assert s in exported, "declared in the header but not exported: " + s
        assert hasattr(lib, s)


def test_ctypes_table_matches_header():
    p = pkg()
    assert sorted(p.capi.PROTOTYPES) == header_symbols()


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = pkg()
    with pytest.raises(p.Cme213Error) as e:
        p.Context(0)
    assert e.value.code in (p.capi.ERR_CUDA, p.capi.ERR_INVALID)
    # host-pointer entry points report the failure too (status, never exit())
    import numpy as np
    x = np.arange(8, dtype=np.uint32)
    rc = p.load().cme213_exclusive_scan_u32_host(p.capi.ptr(x), p.capi.ptr(x.copy()), 8)
    assert rc != 0 and p.load().cme213_last_error()


def test_product_package_never_touches_the_oracle():
    pkg_dir = os.path.join(ROOT, "2012-04_stanford_cme213_b200")
    for root, _, files in os.walk(pkg_dir):
        if os.path.basename(root) == "build":
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(root, f)).read()
                assert not re.search(r"^\s*(import|from)\s+oracle|liboracle|oracle/_ref|#include\s+\"[^\"]*oracle", text, re.M), f
