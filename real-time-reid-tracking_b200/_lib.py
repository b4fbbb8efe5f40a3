"""ctypes loader for libreidb200.so (the C ABI in include/reidb200.h).

There is no CPU fallback: if the library is missing it is built with nvcc; if that is impossible, or no
sm_100 device is visible when a kernel is requested, the call raises.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libreidb200.so")
ABI_VERSION = 8

OK, EINVAL, ECUDA, ELIMIT = 0, -1, -2, -3
RESIZE_CV_LINEAR, RESIZE_PIL_ANTIALIAS = 0, 1
DIST_EUCLID, DIST_COSINE_HALF, DIST_DOT, DIST_SQL2 = 0, 1, 2, 3
IOU_PLAIN, IOU_DIOU = 0, 1
BOX_MODES = {"iou": 0, "giou": 1, "diou": 2, "ciou": 3, "plus_one": 4, "iou_distance": 5}
ASSOC_CLAMP_APP, ASSOC_CLAMP_IOU, ASSOC_NO_GATE, ASSOC_SIMT = 1, 2, 4, 8

_vp, _i, _i64, _sz, _f, _d = C.c_void_p, C.c_int, C.c_int64, C.c_size_t, C.c_float, C.c_double
_f3 = C.POINTER(C.c_float)

# name -> (restype, argtypes); mirrors include/reidb200.h line by line
SIGNATURES = {
    "reidb200_last_error": (C.c_char_p, []),
    "reidb200_abi_version": (_i, []),
    "reidb200_check_device": (_i, []),
    "reidb200_profile_enable": (_i, [_i]),
    "reidb200_profile_read": (_i, [C.c_char_p, C.POINTER(_i), C.POINTER(_d)]),
    "reidb200_profile_names": (_i, [C.c_char_p, _i]),
    "reidb200_crop_resize_norm": (_i, [_vp, _vp, _vp, _i, _i, _i, _f3, _f3, _i, _vp, _vp]),
    "reidb200_crop_resize_norm_ws": (_i, [_vp, _vp, _vp, _i, _i, _i, _f3, _f3, _i, _vp, _vp, _sz, _vp]),
    "reidb200_resize_pil_u8": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, _vp, _sz, _vp]),
    "reidb200_resize_pil_workspace_bytes": (_sz, [_i, _i, _i]),
    "reidb200_distmat_workspace_bytes": (_sz, [_i, _i, _i]),
    "reidb200_distmat": (_i, [_vp, _i, _i64, _vp, _i, _i64, _i, _i, _vp, _i64, _vp, _sz, _vp]),
    "reidb200_distmat_simt": (_i, [_vp, _i, _i64, _vp, _i, _i64, _i, _i, _vp, _i64, _vp]),
    "reidb200_pack_bytes": (_sz, [_i, _i]),
    "reidb200_pack_rows": (_i, [_vp, _i, _i64, _i, _vp, _vp, _vp, _vp]),
    "reidb200_knn_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "reidb200_knn": (_i, [_vp, _i, _i64, _vp, _i, _i64, _i, _i, _i, _i64, _vp, _vp, _i, _vp, _sz, _vp]),
    "reidb200_knn_packed_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "reidb200_knn_packed": (_i, [_vp, _i, _i64, _vp, _vp, _vp, _i, _i, _i, _i, _i64, _vp, _vp, _i, _vp, _sz, _vp]),
    "reidb200_topk_rows": (_i, [_vp, _i, _i, _i64, _vp, _i, _i, _i64, _vp, _vp, _i, _vp]),
    "reidb200_eval_workspace_bytes": (_sz, [_i, _i]),
    "reidb200_eval_cmc_map": (_i, [_vp, _i, _i, _i64, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "reidb200_eval_cmc_map_static": (_i, [_vp, _i, _i, _i64, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "reidb200_jaccard_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "reidb200_jaccard_region_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "reidb200_jaccard": (_i, [_vp, _i, _i64, _i, _i, _i, _i, _f, _i, _i, _i, _i, _vp, _i64, _vp, _vp, _sz, _vp]),
    "reidb200_jaccard_from_rank": (_i, [_vp, _vp, _i64, _vp, _i, _i, _i, _i, _f, _i, _i, _i, _i, _vp, _i64, _vp, _sz, _vp]),
    "reidb200_rerank_rows_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "reidb200_rerank_rows_from_rank": (_i, [_vp, _i, _i64, _i, _vp, _vp, _i, _i, _i, _f, _i, _i, _i, _i, _vp, _i64, _vp, _i64,
                                            _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "reidb200_k_reciprocal": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp]),
    "reidb200_gallery_create": (_i, [_i, _i, _i, C.POINTER(_vp)]),
    "reidb200_gallery_destroy": (_i, [_vp]),
    "reidb200_gallery_resize": (_i, [_vp, _i, _i, _vp]),
    "reidb200_gallery_partial_fit": (_i, [_vp, _vp, _i64, _vp, _i, _vp]),
    "reidb200_gallery_reset_slots": (_i, [_vp, _vp, _i, _vp]),
    "reidb200_gallery_counts": (_i, [_vp, _vp, _vp]),
    "reidb200_gallery_read_slot": (_i, [_vp, _i, _vp, _vp]),
    "reidb200_assoc_frame": (_i, [_vp, _vp, _i64, _vp, _i, _vp, _vp, _vp, _vp, _i, _d, _d, _d, _i, _i, _vp, _vp, _vp, _vp]),
    "reidb200_frame_create": (_i, [_vp, _i, _i, C.POINTER(_vp)]),
    "reidb200_frame_destroy": (_i, [_vp]),
    "reidb200_frame_crops": (_i, [_vp, _vp, _i, _i, _vp, _i, _i, _i, _f3, _f3, _vp, _vp]),
    "reidb200_frame_costs": (_i, [_vp, _vp, _i64, _vp, _i, _vp, _vp, _vp, _vp, _i, _d, _d, _d, _i, _i, C.POINTER(_vp), C.POINTER(_vp), _vp]),
    "reidb200_iou_matrix": (_i, [_vp, _i, _vp, _i, _i, _i, _vp, _vp]),
    "reidb200_gating_distance": (_i, [_vp, _vp, _vp, _i, _i, _vp, _vp]),
    "reidb200_cosine_distance_workspace_bytes": (_sz, [_i, _i, _i]),
    "reidb200_cosine_distance": (_i, [_vp, _i, _i64, _vp, _i, _i64, _i, _i, _vp, _vp, _sz, _vp]),
    "reidb200_rerank_eval_workspace_bytes": (_sz, [_i, _i, _i, _i, _i, _i]),
    "reidb200_rerank_eval": (_i, [_vp, _i, _i, _i64, _i, _i, _i, _f, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _sz, _vp]),
    "reidb200_rerank_eval_pq": (_i, [_vp, _i, _i, _i64, _i, _i, _i, _f, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "reidb200_box_pairs": (_i, [_vp, _i, _vp, _i, _i, _i, _vp, _vp]),
    "reidb200_fuse_cost": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "reidb200_ocsort_cost": (_i, [_vp, _i, _vp, _i, _vp, _vp, _d, _vp, _vp, _vp]),
    "reidb200_kf_initiate": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _d, _d, _vp]),
    "reidb200_kf_predict": (_i, [_vp, _vp, _vp, _i, _d, _d, _d, _vp]),
    "reidb200_kf_update": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _d, _vp]),
    "reidb200_tta_merge_normalize": (_i, [_vp, _sz, _vp, _sz, _i, _i, _vp, _sz, _vp]),
    "reidb200_clamp_add": (_i, [_vp, _i, _i, _sz, _vp, _sz, _vp]),
    "reidb200_dbscan_workspace_bytes": (_sz, [_i]),
    "reidb200_dbscan_precomputed": (_i, [_vp, _i, _sz, _f, _i, _vp, _vp, _vp, _vp, _sz, _vp]),
    "reidb200_smooth_groups": (_i, [_vp, _i, _sz, _vp, _vp, _i, _f, _f, _vp]),
    "reidb200_center_columns": (_i, [_vp, _i, _i, _sz, _vp, _sz, _vp, _vp]),
    "reidb200_lap_workspace_bytes": (_sz, [_i, _i]),
    "reidb200_lap_gated": (_i, [_vp, _i, _i, _sz, _vp, _i, _d, _vp, _vp, _vp, _vp, _sz, _vp]),
    "reidb200_lap_exact": (_i, [_vp, _i, _i, _sz, _vp, _i, _vp, _vp, _i, _d, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
}

_lib = None
_lock = threading.Lock()
_device_ok = set()


class Reidb200Error(RuntimeError):
    pass


def load(build_if_missing: bool = True) -> C.CDLL:
    """Load (building first if needed) libreidb200.so and bind every symbol of the header."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            if not build_if_missing:
                raise Reidb200Error("libreidb200.so is missing (%s): run `python __graft_entry__.py build`" % LIB_PATH)
            from . import build as _build

            _build.build()
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError here == header/library mismatch: fail loudly
            fn.restype = res
            fn.argtypes = args
        if lib.reidb200_abi_version() != ABI_VERSION:
            raise Reidb200Error("libreidb200.so ABI %d != loader ABI %d: rebuild" % (lib.reidb200_abi_version(), ABI_VERSION))
        _lib = lib
    return _lib


def last_error() -> str:
    return load().reidb200_last_error().decode("utf-8", "replace")


def check(rc: int, exc=None):
    """Convert a C status into the exception type the reference would raise."""
    if rc == OK:
        return
    msg = last_error()
    if rc == EINVAL:
        raise (exc or ValueError)(msg)
    if rc == ELIMIT:
        raise NotImplementedError(msg)
    raise Reidb200Error(msg)


def require_device(t: torch.Tensor = None) -> torch.device:
    """The CUDA device work will run on; raises when there is none (no CPU fallback)."""
    if not torch.cuda.is_available():
        raise Reidb200Error("reidb200 needs an sm_100 (B200) CUDA device; none is visible and there is no CPU fallback")
    dev = t.device if (t is not None and t.is_cuda) else torch.device("cuda", torch.cuda.current_device())
    if dev.index not in _device_ok:
        with torch.cuda.device(dev):
            check(load().reidb200_check_device())
        _device_ok.add(dev.index)
    return dev


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream_ptr() -> int:
    """cudaStream_t of torch's current stream on the current device (the raw getter skips building a Stream object)."""
    if _raw_stream is not None:
        return _raw_stream(torch.cuda.current_device())
    return torch.cuda.current_stream().cuda_stream


def ptr(t) -> int:
    return 0 if t is None else t.data_ptr()


def workspace(nbytes: int, device) -> torch.Tensor:
    """256-byte aligned scratch (torch's caching allocator hands out 512-byte aligned blocks)."""
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)


class HostStaging:
    """Small per-call host arrays -> device in ONE copy: the arrays are laid out back to back (8-byte aligned) in a
    pinned buffer, copied with a single asynchronous H2D, and returned as typed views of the device mirror.  Two
    buffer pairs alternate, so the views handed out by one call stay valid until the call after the next (a frame's
    inputs are consumed by the kernels launched right after).  An event per pair guards the pinned memory against
    being rewritten while its previous copy is still in flight."""

    def __init__(self, device):
        self.dev = device
        self.slot = 0
        self.pairs = [dict(cap=0, host=None, dev=None, ev=None) for _ in range(2)]

    def upload(self, arrays):
        import numpy as np

        sizes = [(a.nbytes + 7) & ~7 for a in arrays]
        total = max(sum(sizes), 8)
        pr = self.pairs[self.slot]
        self.slot ^= 1
        if pr["cap"] < total:
            cap = max(4096, 1 << (total - 1).bit_length())
            pr["host"] = torch.empty(cap, dtype=torch.uint8).pin_memory()
            pr["dev"] = torch.empty(cap, dtype=torch.uint8, device=self.dev)
            pr["cap"], pr["ev"], pr["sig"] = cap, None, None
        if pr["ev"] is not None:
            pr["ev"].synchronize()
        sig = tuple((a.dtype.str, a.shape) for a in arrays)
        if pr.get("sig") != sig:  # typed host / device views of this layout, rebuilt only when the shapes change
            hv = pr["host"].numpy()
            off, hviews, dviews = 0, [], []
            for a, sz in zip(arrays, sizes):
                hviews.append(hv[off:off + a.nbytes].view(a.dtype).reshape(a.shape))
                dviews.append(pr["dev"][off:off + a.nbytes].view(_TORCH_OF[a.dtype.str]).reshape(a.shape))
                off += sz
            pr["sig"], pr["hviews"], pr["dviews"], pr["bytes"] = sig, hviews, dviews, off
            pr["src"], pr["dst"] = pr["host"][:off], pr["dev"][:off]
        for h, a in zip(pr["hviews"], arrays):
            h[...] = a
        views, off = pr["dviews"], pr["bytes"]
        pr["dst"].copy_(pr["src"], non_blocking=True)
        if pr["ev"] is None:
            pr["ev"] = torch.cuda.Event()
        pr["ev"].record()
        return views


_TORCH_OF = {"<f8": torch.float64, "<f4": torch.float32, "<i4": torch.int32, "<i8": torch.int64, "|u1": torch.uint8}
_staging = {}


def staging(device) -> HostStaging:
    key = (device.type, device.index)
    if key not in _staging:
        _staging[key] = HostStaging(device)
    return _staging[key]


def as_f32_cuda(x, device=None) -> torch.Tensor:
    """Any array-like -> contiguous float32 CUDA tensor (H2D when it lives on the host)."""
    if not isinstance(x, torch.Tensor):
        import numpy as np

        x = torch.from_numpy(np.ascontiguousarray(np.asarray(x, dtype=np.float32)))
    dev = device or (x.device if x.is_cuda else require_device())
    x = x.to(device=dev, dtype=torch.float32, non_blocking=True)
    return x if x.is_contiguous() else x.contiguous()


def profile_enable(on: bool = True):
    """Bracket every kernel launch of the library with CUDA events on its launch stream (bench.py)."""
    check(load().reidb200_profile_enable(1 if on else 0))


def profile_read() -> dict:
    """{kernel name: (launches, total milliseconds)} of the spans recorded since profile_enable."""
    lib = load()
    buf = C.create_string_buffer(4096)
    check(lib.reidb200_profile_names(buf, 4096))
    out = {}
    for name in buf.value.decode().split("\n"):
        if not name:
            continue
        n, ms = C.c_int(0), C.c_double(0.0)
        check(lib.reidb200_profile_read(name.encode(), C.byref(n), C.byref(ms)))
        out[name] = (n.value, ms.value)
    return out
