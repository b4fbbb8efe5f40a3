"""GPU test of the allocation-free C entry reidb200_eval_cmc_map_static (include/reidb200.h: no counterpart in the reference,
added for callers that cannot cudaMalloc).  Its scratch is a file-scope static __device__ array inside the library that all
calls share and that keeps the previous call's contents, so the test makes TWO calls in a row on different sets (the second
one smaller, so stale histogram counts beyond it would show) and compares each, bit for bit, with the existing
workspace-taking entry reidb200_eval_cmc_map on the same inputs -- same kernels, integer counts and an ordered fp64 sum, so
the tolerance is zero.  Called through the C ABI (ctypes) with every buffer allocated BEFORE the calls; torch's allocator
statistics show that nothing was allocated in between.  Collected last on purpose (file name): the round's end runs -x."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _set(Q, G, D, seed, noise):
    from reidb200 import synth

    d = synth.market_like(Q=Q, G=G, D=D, n_ids=max(G // 12, 4), seed=seed, orphan_frac=0.05, noise=noise)
    dev = torch.device("cuda", 0)
    qf, gf = torch.from_numpy(d["qf"]).to(dev), torch.from_numpy(d["gf"]).to(dev)
    score = (qf @ gf.T).contiguous()  # any score matrix will do: both entries read the same one
    lab = [torch.from_numpy(d[k].astype(np.int32)).to(dev) for k in ("ql", "qc", "gl", "gc")]
    return score, lab


def _outputs(Q, G):
    dev = torch.device("cuda", 0)
    return (torch.full((G,), -7.0, dtype=torch.float32, device=dev), torch.full((1,), -7.0, dtype=torch.float64, device=dev),
            torch.full((Q,), -7.0, dtype=torch.float64, device=dev), torch.full((Q,), -7, dtype=torch.int32, device=dev))


def test_static_scratch_entry_equals_the_workspace_entry_twice_in_a_row(rb):
    from reidb200 import _lib as L

    lib = L.load()
    torch.cuda.set_device(0)
    cases = [(700, 5000, 64, 5, 1.6, 1), (130, 901, 32, 6, 1.2, 0), (700, 5000, 64, 5, 1.6, 0)]
    prepared = []
    for Q, G, D, seed, noise, desc in cases:
        score, lab = _set(Q, G, D, seed, noise)
        if not desc:
            score = (2.0 - 2.0 * score).contiguous()  # a distance matrix for the ascending mode
        ws = torch.empty(lib.reidb200_eval_workspace_bytes(Q, G), dtype=torch.uint8, device=score.device)
        prepared.append((Q, G, desc, score, lab, ws, _outputs(Q, G), _outputs(Q, G), _outputs(Q, G)))
    torch.cuda.synchronize()
    st = L.stream_ptr()
    before = torch.cuda.memory_stats(0)["allocation.all.allocated"]
    for Q, G, desc, score, lab, ws, ref, got, got2 in prepared:
        L.check(lib.reidb200_eval_cmc_map(score.data_ptr(), Q, G, score.stride(0), desc, *[t.data_ptr() for t in lab],
                                          ref[0].data_ptr(), ref[1].data_ptr(), ref[2].data_ptr(), ref[3].data_ptr(),
                                          ws.data_ptr(), ws.numel(), st))
        # with the per-query outputs ...
        L.check(lib.reidb200_eval_cmc_map_static(score.data_ptr(), Q, G, score.stride(0), desc, *[t.data_ptr() for t in lab],
                                                 got[0].data_ptr(), got[1].data_ptr(), got[2].data_ptr(), got[3].data_ptr(), st))
        # ... and without them (per-query AP / first hit then live in the static array too), right behind the first call
        L.check(lib.reidb200_eval_cmc_map_static(score.data_ptr(), Q, G, score.stride(0), desc, *[t.data_ptr() for t in lab],
                                                 got2[0].data_ptr(), got2[1].data_ptr(), None, None, st))
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats(0)["allocation.all.allocated"] == before  # nothing went through torch's allocator
    for Q, G, desc, score, lab, ws, ref, got, got2 in prepared:
        assert 0.05 < float(ref[0][0]) < 0.999 and float(ref[1]) > 0  # a curve, not a constant
        for a, b in zip(ref, got):
            assert torch.equal(a, b)
        assert torch.equal(ref[0], got2[0]) and torch.equal(ref[1], got2[1])
        assert float(got2[2][0]) == -7.0 and int(got2[3][0]) == -7  # untouched when not requested


def test_static_scratch_entry_rejects_sizes_beyond_its_maximum(rb):
    from reidb200 import _lib as L

    lib = L.load()
    torch.cuda.set_device(0)
    score, lab = _set(8, 40, 16, 1, 1.0)
    out = _outputs(8, 40)
    args = [t.data_ptr() for t in lab] + [out[0].data_ptr(), out[1].data_ptr(), None, None, L.stream_ptr()]
    for Q, G in ((8193, 40), (8, 32769)):
        with pytest.raises(ValueError, match="exceed the static scratch"):
            L.check(lib.reidb200_eval_cmc_map_static(score.data_ptr(), Q, G, max(G, 40), 1, *args))
    with pytest.raises(ValueError, match="empty"):
        L.check(lib.reidb200_eval_cmc_map_static(score.data_ptr(), 0, 40, 40, 1, *args))
    torch.cuda.synchronize()
    assert float(out[0][0]) == -7.0  # a rejected call launches nothing
