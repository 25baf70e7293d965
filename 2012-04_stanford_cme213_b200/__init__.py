"""cme213-b200: B200-native CSR/ELL/HYB SpMV, scan and transpose behind a C ABI.

The directory name is not a Python identifier; import it with
    importlib.import_module("2012-04_stanford_cme213_b200")
(see tests/conftest.py and bench.py).  The compute path is libcme213_b200.so; this package only
binds it (capi) and hosts the multi-GPU row-sharding logic (sharded).
"""
from . import capi  # noqa: F401
from .capi import Context, Cme213Error, load  # noqa: F401
