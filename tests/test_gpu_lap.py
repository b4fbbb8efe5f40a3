"""N1: min_cost_matching's assignment on the device against scipy's linear_sum_assignment run the way the
reference runs it (linear_assignment.py:58-77: clamp, solve, drop pairs above max_distance)."""
import numpy as np
import pytest
import torch

from oracle import reid_oracle as O

pytestmark = pytest.mark.gpu


def _reference_matches(cost, max_distance):
    """The oracle's restatement of min_cost_matching (clamp, scipy's solve, post-filter; pinned by assoc_seq.npz)."""
    T, M = cost.shape
    matches, _, _ = O.min_cost_matching_from_cost(cost, max_distance, range(T), range(M))
    return sorted((int(r), int(k)) for r, k in matches)


def _device_matches(cost, max_distance, row_cand=None, match=None, det=None):
    from reidb200 import ops

    T, M = cost.shape
    c = torch.from_numpy(cost).cuda()
    match = torch.full((T,), -1, dtype=torch.int32, device="cuda") if match is None else match
    det = torch.zeros(M, dtype=torch.int32, device="cuda") if det is None else det
    rows = np.arange(T, dtype=np.int32) if row_cand is None else np.asarray(row_cand, np.int32)
    status = ops.lap_gated(c, rows, max_distance, match, det)
    assert int(status.item()) == 0
    m = match.cpu().numpy()
    d = det.cpu().numpy()
    pairs = sorted((int(r), int(m[r])) for r in range(T) if m[r] >= 0)
    assert sorted(k for _, k in pairs) == [j for j in range(M) if d[j]]
    return pairs, match, det


@pytest.mark.parametrize("T,M,admissible", [(100, 30, 0.03), (30, 100, 0.05), (300, 200, 0.01), (200, 300, 0.02), (64, 64, 0.2),
                                            (40, 60, 1.0), (60, 40, 1.0), (300, 200, 1.0), (1, 7, 1.0), (7, 1, 0.5)])
def test_lap_matches_scipy(rb, T, M, admissible):
    """Random float64 costs with a fraction of admissible entries: many small components (tracker frames), one large
    component, and the fully dense matrix; both orientations (scipy transposes the taller problem)."""
    rng = np.random.default_rng(T * 1000 + M)
    max_d = 0.2
    cost = rng.uniform(0.21, 1.0, (T, M))
    keep = rng.random((T, M)) < admissible
    cost[keep] = rng.uniform(0.0, 0.2, int(keep.sum()))
    got, _, _ = _device_matches(cost, max_d)
    assert got == _reference_matches(cost, max_d)
    if admissible < 1.0:
        assert len(got) > 0


def test_lap_chained_stages_and_subsets(rb):
    """Stage A on a subset of rows, stage B on other rows and the columns stage A left, as Tracker._match chains them."""
    rng = np.random.default_rng(9)
    T, M = 120, 90
    a = rng.uniform(0.0, 1.0, (T, M))
    a[rng.random((T, M)) < 0.93] = 5.0
    b = rng.uniform(0.0, 1.0, (T, M))
    b[rng.random((T, M)) < 0.9] = 1.0
    rows_a = [r for r in range(T) if r % 3 != 0]
    rows_b = [r for r in range(T) if r % 3 == 0] + rows_a[:20]   # some stage-A rows again: only the unmatched ones count
    got_a, match, det = _device_matches(a, 0.2, rows_a)
    ref_a = [(rows_a[r], k) for r, k in _reference_matches(a[rows_a], 0.2)]
    assert got_a == sorted(ref_a) and len(ref_a) > 5
    got_ab, match, det = _device_matches(b, 0.7, rows_b, match, det)
    taken_r, taken_c = set(r for r, _ in ref_a), set(k for _, k in ref_a)
    rb_left = [r for r in rows_b if r not in taken_r]
    cols_left = [k for k in range(M) if k not in taken_c]
    ref_b = [(rb_left[r], cols_left[k]) for r, k in _reference_matches(b[np.ix_(rb_left, cols_left)], 0.7)]
    assert got_ab == sorted(ref_a + ref_b) and len(ref_b) > 5


def test_lap_edges(rb):
    from reidb200 import ops

    # nothing admissible: no match; non-finite entries count as gated
    cost = np.full((6, 5), 3.0)
    cost[2, 2] = np.nan
    cost[1, 3] = np.inf
    assert _device_matches(cost, 0.5)[0] == []
    # equal to the threshold is admissible (cost > max_distance is the reference's gate)
    cost[4, 0] = 0.5
    assert _device_matches(cost, 0.5)[0] == [(4, 0)]
    with pytest.raises(ValueError):
        ops.lap_gated(torch.zeros((3, 3), device="cuda"), [0], 0.5, torch.zeros(3, dtype=torch.int32, device="cuda"),
                      torch.zeros(3, dtype=torch.int32, device="cuda"))


def test_tracker_match_on_device_equals_host(rb, golden):
    """Tracker._match with both assignment stages on the device: the recorded reference sequence and an MOT20-shaped
    frame give the host path's matches; unmatched sets are equal."""
    from reidb200 import matching as MT, synth, tracker as TR
    from test_gpu_assoc import _frames, _metric_for

    g = golden("assoc_seq")
    max_dist, max_iou, lam, budget = (float(g["cfg_max_dist"]), float(g["cfg_max_iou_dist"]), float(g["cfg_mc_lambda"]),
                                      int(g["cfg_budget"]))
    n = 0
    for fr in _frames(g):
        metric = _metric_for(fr, budget, max_dist)
        m, ut, ud = TR.match_arrays(metric, fr["det_feat"], fr["det_tlwh"], fr["trk_id"], fr["trk_confirmed"], fr["trk_mean"],
                                    fr["trk_SI"], fr["trk_tsu"], lam, max_iou, max_dist, device_lap=True)
        assert np.array_equal(np.array(sorted(m), np.int64).reshape(-1, 2), fr["matches"])
        assert sorted(ut) == fr["unmatched_tracks"].tolist() and sorted(ud) == fr["unmatched_dets"].tolist()
        n += len(m)
    assert n > 20
    M, T = 200, 300
    fr = synth.mot_frame(M=M, T=T, budget=100, D=512, seed=5)
    metric = MT.NearestNeighborDistanceMetric("cosine", 0.2, 100, max_tracks=T)
    targets = np.repeat(np.arange(T), [len(s) for s in fr["samples"]])
    metric.partial_fit(np.concatenate(fr["samples"], 0), targets, list(range(T)))
    conf = np.arange(T) % 7 != 0
    args = (metric, fr["det_feat"], fr["det_tlwh"], np.arange(T), conf, fr["trk_mean"], fr["trk_SI"], fr["trk_tsu"], 0.995, 0.7, 0.2)
    m_h, ut_h, ud_h = TR.match_arrays(*args, device_lap=False)
    m_d, ut_d, ud_d = TR.match_arrays(*args, device_lap=True)
    assert sorted(m_d) == sorted(m_h) and len(m_h) > 50 and sorted(ut_d) == sorted(ut_h) and sorted(ud_d) == sorted(ud_h)
    # exact mode: the three lists in the reference's ORDER (the order of unmatched_detections decides the ids of new tracks)
    assert m_d == m_h and ut_d == ut_h and ud_d == ud_h
    # rows spread so that CPython's set order of the unmatched confirmed tracks is not ascending (the row order of stage B)
    conf2 = np.ones(T, bool)
    conf2[::5] = False
    tsu2 = np.where(np.arange(T) % 3 == 0, 1, fr["trk_tsu"]).astype(np.int32)
    args2 = (metric, fr["det_feat"], fr["det_tlwh"][:40], np.arange(T), conf2, fr["trk_mean"], fr["trk_SI"], tsu2, 0.995, 0.7, 0.2)
    args2 = (metric, fr["det_feat"][:40], fr["det_tlwh"][:40]) + args2[3:]
    h2, d2 = TR.match_arrays(*args2, device_lap=False), TR.match_arrays(*args2, device_lap=True)
    assert d2[0] == h2[0] and d2[1] == h2[1] and d2[2] == h2[2] and len(h2[1]) > 100
    m_f, ut_f, ud_f = TR.match_arrays(*args, device_lap="fast")
    assert sorted(m_f) == sorted(m_h) and sorted(ut_f) == sorted(ut_h) and ud_f == sorted(ud_h)


def _exact_on_device(cost, max_distance, rows, cols):
    from reidb200 import ops

    T, M = cost.shape
    dev = torch.device("cuda", 0)
    c = torch.from_numpy(cost).to(dev)
    match = torch.full((T,), -1, dtype=torch.int32, device=dev)
    det = torch.zeros(M, dtype=torch.int32, device=dev)
    um = torch.full((max(len(cols), 1),), -7, dtype=torch.int32, device=dev)
    n = torch.zeros(1, dtype=torch.int32, device=dev)
    rows_d = torch.tensor(rows, dtype=torch.int32, device=dev)
    cols_d = torch.tensor(cols, dtype=torch.int32, device=dev)
    status = ops.lap_exact(c, rows_d, cols_d, None, max_distance, match, det, um, n)
    assert int(status.item()) == 0
    m = match.cpu().numpy()
    matches = [(r, int(m[r])) for r in rows if m[r] >= 0]
    return matches, um.cpu().numpy()[:int(n.item())].tolist(), det.cpu().numpy()


@pytest.mark.parametrize("T,M", [(40, 25), (25, 40), (33, 33), (300, 200), (200, 300), (1, 9), (9, 1), (70, 64)])
def test_lap_exact_mode_reproduces_scipys_ties_and_list_order(rb, T, M):
    """Exact mode against scipy run the reference's way on TIE-HEAVY matrices (gated entries all equal, coarse costs, rows
    without any admissible entry): matches in row order and the unmatched_detections list in the reference's order; on
    candidate subsets given in shuffled order too (stage B of Tracker._match gets its columns in stage A's list order)."""
    rng = np.random.default_rng(T * 1000 + M)
    for trial in range(6):
        cost = rng.random((T, M))
        thr = [0.02, 0.1, 0.4, 0.1, 0.1, 5.0][trial]
        if trial == 3:
            cost = np.round(cost, 1)           # exact ties between admissible entries as well
        if trial == 4:
            cost[:] = 0.9                      # nothing admissible: the pairing is pure tie-breaking
        rows, cols = list(range(T)), list(range(M))
        if trial in (1, 2) and min(T, M) > 4:
            rows = rng.permutation(T)[: max(2, T * 2 // 3)].tolist()
            cols = rng.permutation(M)[: max(2, M * 3 // 4)].tolist()
        sub = cost[np.ix_(rows, cols)]
        want_m, want_ut, want_ud = O.min_cost_matching_lists(sub, thr, rows, cols)
        got_m, got_ud, det = _exact_on_device(cost, thr, rows, cols)
        assert got_m == [(int(a), int(b)) for a, b in want_m], (T, M, trial)
        assert got_ud == [int(x) for x in want_ud], (T, M, trial)
        assert sorted(np.nonzero(det)[0].tolist()) == sorted(int(b) for _, b in want_m)
        if trial == 4:
            assert not want_m and len(want_ud) == len(cols)
