"""Host-side model of the triangle-mode selection rule (csrc/gemm.cu GemmArgs::lower + csrc/topk.cu TRI), no GPU: for every
row, the candidate set the row select builds -- direct chunks read in full, chunks of the half that is not stored taken as
(minimum, place) when the packed lower bound of their SECOND smallest key cannot pass, gathered in full otherwise -- must
equal {c : key(r, c) <= t0}, with t0 the k-th smallest of 64 interleaved group minima, ties and duplicates included.
The CUDA path is compared with the full-matrix mode bit for bit in tests/test_gpu_zy_triangle.py; this test pins the rule
itself (the 5 low bits of the second key are replaced by the place, which must only ever make the test conservative)."""
import numpy as np
import pytest


def _ordered(f):
    u = np.asarray(f, np.float32).view(np.uint32)
    return np.where(u & np.uint32(0x80000000), ~u, u | np.uint32(0x80000000)).astype(np.uint32)


@pytest.mark.parametrize("N,k,dup", [(1024, 20, 40), (2080, 12, 0), (1312, 32, 300)])
def test_single_or_gather_rule_reproduces_the_threshold_set(N, k, dup):
    rng = np.random.default_rng(N)
    x = rng.standard_normal((N, 8)).astype(np.float32)
    if dup:
        x[100:100 + dup] = x[100]  # exact ties: second == minimum inside a chunk must force the gather
    d = ((x[:, None, :] - x[None, :, :]) ** 2).sum(-1).astype(np.float32)
    d = np.minimum(d, d.T)  # what the matrix is: symmetric bit for bit
    key = _ordered(d)
    nch = (N + 31) // 32
    pad = np.full((N, nch * 32), 0xFFFFFFFF, np.uint32)
    pad[:, :N] = key
    chunks = pad.reshape(N, nch, 32)
    cmin = chunks.min(2)
    part = np.partition(chunks, 1, axis=2)
    second = part[:, :, 1]
    arg = chunks.argmin(2)  # lowest place among equals, like the epilogue's descending select chain
    word2 = (second & np.uint32(0xFFFFFFE0)) | arg.astype(np.uint32)
    for r in range(0, N, 5):
        rt = r >> 8
        groups = np.full(64, 0xFFFFFFFF, np.uint32)
        for c in range(nch):
            groups[c % 64] = min(groups[c % 64], cmin[r, c])
        t0 = np.sort(groups)[min(k, N) - 1]
        cand = []
        for c in np.nonzero(cmin[r] <= t0)[0]:
            if (c >> 3) > rt and (word2[r, c] & np.uint32(0xFFFFFFE0)) > t0:
                cand.append(c * 32 + int(word2[r, c] & 31))
                continue
            cand.extend((c * 32 + np.nonzero(chunks[r, c] <= t0)[0]).tolist())
        truth = np.nonzero(key[r] <= t0)[0]
        assert sorted(cand) == truth.tolist()
        assert len(truth) >= min(k, N)
        # stored-once rule: an element and its transposed place are never both missing, and the diagonal blocks are whole
        c = rng.integers(0, N)
        assert ((r >> 8) >= (c >> 8)) or ((c >> 8) >= (r >> 8))
