"""reidb200 -- B200-native ReID matching hot path (drop-in for the reference's
evaluate / faiss_utils / losses.utils / boxmot matching + association / feature_extractor
preprocessing functions).  All compute runs in hand-written sm_100a CUDA kernels behind the
C ABI declared in include/reidb200.h; there is no CPU fallback."""
from . import synth  # noqa: F401

__all__ = ["synth"]
