"""CPU: host-side logic of the product package and the C ABI surface (no kernel is launched)."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "reidb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(reidb200_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_header_symbol(rb):
    from reidb200 import _lib

    lib = _lib.load()
    names = _header_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), "libreidb200.so lacks %s declared in include/reidb200.h" % n
    assert set(names) == set(_lib.SIGNATURES), set(names) ^ set(_lib.SIGNATURES)
    assert lib.reidb200_abi_version() == _lib.ABI_VERSION


def test_workspace_queries_are_pure(rb):
    from reidb200 import _lib

    lib = _lib.load()
    assert lib.reidb200_distmat_workspace_bytes(0, 5, 5) == 0
    a = lib.reidb200_distmat_workspace_bytes(3368, 15913, 512)
    assert a >= (3368 + 15913) * 2 * 512 * 2 and a % 256 == 0
    assert lib.reidb200_jaccard_workspace_bytes(19281, 512, 20, 6) > 19281 * 19281 * 4
    assert lib.reidb200_knn_workspace_bytes(10, 100, 64, 5) > 0
    assert lib.reidb200_pack_bytes(10, 100) == 10 * 2 * 128 * 2 + (256 - (10 * 2 * 128 * 2) % 256) % 256


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback(rb):
    from reidb200 import _lib, faiss_utils, losses_utils

    with pytest.raises(_lib.Reidb200Error):
        losses_utils.euclidean_dist(torch.zeros(2, 4), torch.zeros(3, 4))
    with pytest.raises(_lib.Reidb200Error):
        faiss_utils.compute_jaccard_distance(torch.zeros(30, 4), print_flag=False)


def test_xywh_to_xyxy(rb, golden):
    from reidb200 import preprocess as P

    g = golden("preprocess")
    H, W = g["img"].shape[:2]
    assert np.array_equal(P.xywh_to_xyxy(g["boxes_xywh"], W, H), g["xyxy"])
    assert np.array_equal(P.xywh_to_xyxy(torch.from_numpy(g["boxes_xywh"]), W, H), g["xyxy"])
    # float32 arithmetic like the reference's per-box code on a float32 tensor row
    b = np.array([[100.49999, 50.25, 20.999998, 31.0]], np.float32)
    x, y, w, h = b[0]
    want = [max(int(x - w / 2), 0), max(int(y - h / 2), 0), min(int(x + w / 2), W - 1), min(int(y + h / 2), H - 1)]
    assert P.xywh_to_xyxy(b, W, H)[0].tolist() == want


def test_faiss_shim_assertions(rb):
    from reidb200 import faiss_utils as F

    x = torch.zeros(4, 6)
    assert F.swig_ptr_from_FloatTensor(x) == x.data_ptr()
    with pytest.raises(AssertionError):
        F.swig_ptr_from_FloatTensor(x.t())
    with pytest.raises(AssertionError):
        F.swig_ptr_from_LongTensor(x)
    with pytest.raises(AssertionError):
        F.swig_ptr_from_IntTensor(x)
    with pytest.raises(TypeError):
        F.search_raw_array_pytorch(None, torch.zeros(8, 6), torch.zeros(4, 12)[:, ::2], 2)
