"""Comparison rules of BASELINE.json's north_star: ids identical wherever adjacent distances differ by more than
1e-5 relative, distances within 1e-5 relative."""
import numpy as np

REL_TOL = 1e-5  # north_star: "distances within 1e-5 relative"


def rel_close(a, b, tol=REL_TOL):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return np.abs(a - b) <= tol * np.maximum(np.maximum(np.abs(a), np.abs(b)), 1e-30) + 1e-12


def check_parity(ids_g, dist_g, cnt_g, ids_o, dist_o, cnt_o, tol=REL_TOL, what=""):
    """Asserts the parity bar. Returns (number of id positions that differ inside a distance tie, total)."""
    ids_g, ids_o = np.asarray(ids_g), np.asarray(ids_o)
    assert ids_g.shape == ids_o.shape, what
    assert np.array_equal(np.asarray(cnt_g, dtype=np.int64), np.asarray(cnt_o, dtype=np.int64)), \
        "%s: result counts differ" % what
    valid = ids_o >= 0
    assert np.array_equal(ids_g >= 0, valid), "%s: padding differs" % what
    ok_d = rel_close(dist_g, dist_o, tol) | ~valid
    bad = np.argwhere(~ok_d)
    assert bad.size == 0, "%s: %d distances differ by more than %g relative, first at %s: %r vs %r" % (
        what, len(bad), tol, bad[0], dist_g[tuple(bad[0])], dist_o[tuple(bad[0])])
    diff = (ids_g != ids_o) & valid
    # a differing id is acceptable only inside a tie: same position, distances equal to tolerance (checked above),
    # and the oracle's own neighbours at that rank are within tolerance of each other (or it is the k-th boundary)
    ties = 0
    for q, j in np.argwhere(diff):
        d = float(dist_o[q, j])
        near = False
        for jj in (j - 1, j + 1):
            if 0 <= jj < ids_o.shape[1] and ids_o[q, jj] >= 0 and rel_close(dist_o[q, jj], d, tol):
                near = True
        if j == ids_o.shape[1] - 1 or j == int(cnt_o[q]) - 1:
            near = True  # boundary: the tied partner may be the (k+1)-th candidate, which is not reported
        assert near, "%s: query %d rank %d: id %d vs oracle %d with no distance tie (d=%r)" % (
            what, q, j, ids_g[q, j], ids_o[q, j], d)
        ties += 1
    return ties, int(valid.sum())


def parity_report(ids_g, dist_g, ids_o, dist_o, cnt_o, tol=REL_TOL):
    """The same rules as check_parity() as COUNTS instead of assertions (bench.py prints them for the whole batch).
    A differing id is "inside a tie" when the distance at that rank agrees to `tol` and the oracle's own neighbour at
    rank j-1 / j+1 is within `tol` of it (or the rank is the k-th boundary); everything else is "outside ties"."""
    ids_g, ids_o = np.asarray(ids_g), np.asarray(ids_o)
    dist_g, dist_o = np.asarray(dist_g), np.asarray(dist_o)
    valid = (ids_o >= 0) & (ids_g >= 0)
    same = ids_g == ids_o
    diff = ~same & ((ids_o >= 0) | (ids_g >= 0))
    denom = np.maximum(np.maximum(np.abs(dist_g), np.abs(dist_o)).astype(np.float64), 1e-30)
    rel = np.abs(dist_g.astype(np.float64) - dist_o.astype(np.float64)) / denom
    rel_same = rel[same & valid]
    inside = outside = 0
    bad_q = set()
    for q, j in np.argwhere(diff):
        ok = bool(valid[q, j]) and bool(rel_close(dist_g[q, j], dist_o[q, j], tol))
        if ok:
            near = j == ids_o.shape[1] - 1 or j == int(cnt_o[q]) - 1
            d = float(dist_o[q, j])
            for jj in (j - 1, j + 1):
                if 0 <= jj < ids_o.shape[1] and ids_o[q, jj] >= 0 and rel_close(dist_o[q, jj], d, tol):
                    near = True
            ok = near
        if ok:
            inside += 1
        else:
            outside += 1
            bad_q.add(int(q))
    return {"queries": int(ids_o.shape[0]),
            "id_identical_queries": int(same.all(axis=1).sum()),
            "id_mismatch_inside_ties": inside,
            "id_mismatch_outside_ties": outside,
            "queries_with_mismatch_outside_ties": len(bad_q),
            "dist_max_rel_err": float(rel_same.max()) if rel_same.size else 0.0}
