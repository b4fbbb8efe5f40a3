"""Tie-aware comparison of a ranking-based evaluation with the reference's (test helper, host numpy).

north_star: "neighbour and rank indices identical except for ties within that tolerance [distances within 1e-4 relative],
CMC/mAP within 1e-6 absolute".  For a query whose good items are separated from their neighbours in the ranking by more
than the distance tolerance, the AP and the first-hit rank must be THE reference's values; where a good item and a
non-good item are closer than the tolerance either order is a legitimate outcome, and the reference's AP must lie between
the best and the worst arrangement of every such tie group.  `ap_bounds` computes that interval from a distance row."""
import numpy as np


def good_junk(ql, qc, gl, gc):
    """Boolean masks (good, junk) over the gallery for one query (reid/evaluate.py:66-72)."""
    same_id = gl == ql
    same_cam = gc == qc
    good = same_id & ~same_cam
    junk = (same_id & same_cam) | (gl == -1)
    return good, junk


def _ap_from_rows(rows_good):
    """reid/evaluate.py:94-103 on 0-based junk-free positions of the good items (ascending)."""
    n = len(rows_good)
    ap = 0.0
    for i, r in enumerate(rows_good):
        precision = (i + 1) * 1.0 / (r + 1)
        old_precision = i * 1.0 / r if r != 0 else 1.0
        ap += (1.0 / n) * (old_precision + precision) / 2
    return ap


def ap_bounds(dist_row, good, junk, rel_tol=2e-6, abs_tol=0.0):
    """(ap_lo, ap_hi, first_lo, first_hi) over every ordering that agrees with `dist_row` (ascending = better) up to
    rel_tol * |d| + abs_tol: consecutive junk-free entries closer than that are chained into one tie group; inside a
    group the good items take the first (best case) or the last (worst case) places.  No good item: (0, 0, -1, -1)."""
    keep = ~junk
    d = np.asarray(dist_row, np.float64)[keep]
    g = np.asarray(good)[keep]
    if not g.any():
        return 0.0, 0.0, -1, -1
    order = np.argsort(d, kind="stable")
    d, g = d[order], g[order]
    gap_ok = (d[1:] - d[:-1]) <= rel_tol * np.abs(d[1:]) + abs_tol
    gid = np.concatenate([[0], np.cumsum(~gap_ok)])  # tie-group id per position
    starts = np.flatnonzero(np.concatenate([[True], ~gap_ok]))
    sizes = np.diff(np.concatenate([starts, [len(d)]]))
    ngood = np.bincount(gid, weights=g, minlength=len(starts)).astype(np.int64)
    best, worst = [], []
    for s, sz, ng in zip(starts, sizes, ngood):
        if ng:
            best.extend(range(s, s + ng))
            worst.extend(range(s + sz - ng, s + sz))
    return _ap_from_rows(worst), _ap_from_rows(best), best[0], worst[0]


def check_per_query(dist, ql, qc, gl, gc, ap_ref, first_ref, ap_got=None, first_got=None, rel_tol=2e-6, abs_tol=0.0,
                    rows=None):
    """Assert, query by query, that the reference's AP / first hit (and, if given, the kernel's own per-query outputs)
    are consistent with the distance rows `dist` (host f32[Q', G], rows = query ids of those rows or None = all).
    Returns (number of queries with a non-degenerate interval, sum of interval widths)."""
    gl, gc = np.asarray(gl), np.asarray(gc)
    loose, width = 0, 0.0
    ids = range(dist.shape[0]) if rows is None else rows
    for r, q in enumerate(ids):
        good, junk = good_junk(ql[q], qc[q], gl, gc)
        lo, hi, f_lo, f_hi = ap_bounds(dist[r], good, junk, rel_tol, abs_tol)
        eps = 1e-9
        assert lo - eps <= ap_ref[q] <= hi + eps, "query %d: reference AP %.12f outside [%.12f, %.12f]" % (q, ap_ref[q], lo, hi)
        assert f_lo <= first_ref[q] <= f_hi, "query %d: reference first hit %d outside [%d, %d]" % (q, first_ref[q], f_lo, f_hi)
        if ap_got is not None:
            assert lo - eps <= ap_got[q] <= hi + eps, "query %d: kernel AP %.12f outside [%.12f, %.12f]" % (q, ap_got[q], lo, hi)
        if first_got is not None:
            assert f_lo <= first_got[q] <= f_hi, "query %d: kernel first hit %d outside [%d, %d]" % (q, first_got[q], f_lo, f_hi)
        if hi - lo > 0 or f_hi != f_lo:
            loose += 1
            width += hi - lo
    return loose, width
