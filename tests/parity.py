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
