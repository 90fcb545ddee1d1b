"""pangenome-index_b200 — B200-native find_mems hot path (MEM finding over the r-index + tag lookup).

Python side = thin ctypes plumbing over the C ABI in include/pgi.h (libpgi.so, built from csrc/ by
`make -C pangenome-index_b200/csrc`) plus the CPU data-prep tools (prep/). There is no Python or CPU
implementation of the query path: `capi.lib()` raises if the CUDA library is missing.

The directory name contains a hyphen, so import it through `__graft_entry__.load_package()` (or
tests/conftest.py), which registers it as the module `pangenome_index_b200`.
"""
from . import capi, prep, sharding  # noqa: F401

def __getattr__(name):
    # synth / pangraph import torch (data prep for the synthetic workloads); loaded on first use only
    if name in ("synth", "pangraph"):
        import importlib
        return importlib.import_module("." + name, __name__)
    raise AttributeError(name)
