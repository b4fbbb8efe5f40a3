"""Import the reference's own Python modules in place from /root/reference.

TEST INFRASTRUCTURE ONLY, and only usable in the build container (the GPU box has
no /root/reference).  Used by oracle/make_golden.py to pin the oracle and to write
the fixtures under tests/golden/.

Third-party modules the reference imports at module scope but that are absent here
are replaced by minimal stand-ins:
  faiss     exact fp32 squared-L2 brute force (IndexFlatL2 / get_num_gpus / METRIC_L2)
            -- faiss-gpu 1.7.4 (README.md:42-43), "parity unpinned" at this boundary
  lap       never executed on the StrongSORT path
  filterpy  only common.reshape_z is executed (filterpy>=1.4.5 semantics)
  gdown     download helper, never executed
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np

REF = "/root/reference"


def available() -> bool:
    return os.path.isdir(os.path.join(REF, "reid"))


def _faiss_stub():
    from oracle import reid_oracle as O

    m = types.ModuleType("faiss")
    m.METRIC_L2 = 1

    class IndexFlatL2:
        def __init__(self, d):
            self.d = d
            self.xb = np.zeros((0, d), np.float32)

        def add(self, x):
            self.xb = np.concatenate([self.xb, np.asarray(x, np.float32)], 0)

        def reset(self):
            self.xb = np.zeros((0, self.d), np.float32)

        def search(self, xq, k):
            return O.knn_l2(self.xb, np.asarray(xq, np.float32), k)

    m.IndexFlatL2 = IndexFlatL2
    m.get_num_gpus = lambda: 0
    return m


def _filterpy_stub():
    fp = types.ModuleType("filterpy")
    common = types.ModuleType("filterpy.common")
    stats = types.ModuleType("filterpy.stats")

    def reshape_z(z, dim_z, ndim):
        z = np.atleast_2d(z)
        if z.shape[1] == dim_z:
            z = z.T
        if z.shape != (dim_z, 1):
            raise ValueError("z must be convertible to shape ({}, 1)".format(dim_z))
        if ndim == 1:
            z = z[:, 0]
        if ndim == 0:
            z = z[0, 0]
        return z

    common.reshape_z = reshape_z
    common.pretty_str = lambda label, arr: "{} = {}".format(label, arr)
    stats.logpdf = lambda *a, **k: 0.0
    fp.common = common
    fp.stats = stats
    return {"filterpy": fp, "filterpy.common": common, "filterpy.stats": stats}


_installed = False


def install():
    """Put the stubs and the reference roots on sys.modules / sys.path (idempotent)."""
    global _installed
    if _installed:
        return
    if not available():
        raise RuntimeError("/root/reference is not present on this machine")
    sys.modules.setdefault("faiss", _faiss_stub())
    sys.modules.setdefault("lap", types.ModuleType("lap"))
    sys.modules.setdefault("gdown", types.ModuleType("gdown"))
    for k, v in _filterpy_stub().items():
        sys.modules.setdefault(k, v)
    for p in (REF, os.path.join(REF, "yolov8_tracking")):
        if p not in sys.path:
            sys.path.insert(0, p)
    _installed = True


def load_file(name: str, relpath: str):
    """Load one reference file as a stand-alone module (no package side effects)."""
    import importlib.util

    install()
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, relpath))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def load_file_pkg(dotted: str):
    """Import a boxmot submodule by dotted name without executing boxmot/__init__.py
    (which pulls in every tracker and their weights helpers): a bare namespace stand-in
    for the top-level package is registered first."""
    import importlib

    install()
    if "boxmot" not in sys.modules:
        pkg = types.ModuleType("boxmot")
        pkg.__path__ = [os.path.join(REF, "yolov8_tracking", "boxmot")]
        sys.modules["boxmot"] = pkg
    return importlib.import_module(dotted)
