"""TEST INFRASTRUCTURE ONLY - the north_star parity rule for search results, stated as code.

BASELINE.json north_star: "Returned indices are identical except where PQ distances tie within 1e-5 relative.
Recall@K overlap with the reference is at least 0.99 at the same EF/K."

A PQ distance is a function of the vertex (dis(v) = sum_m T[m][code[v][m]]), so two result rows of the same query agree
under that rule iff
  (a) their SORTED distance vectors agree element-wise within rtol (same multiset of distances up to rounding), and
  (b) the ids present in one row and absent from the other pair up one-to-one with partners of equal distance (within
      rtol) - i.e. ids move only inside an equal-distance class, in particular across the k-th position of a class cut
      by the end of the row.
Order inside a row is then free only inside an equal-distance class by construction (rows are ascending in distance).
`tie_rule_report` counts the rows that break (a) or (b); the tests require 0.
"""
from __future__ import annotations

import numpy as np

RTOL = 1e-5


def _close(a, b, rtol):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return np.abs(a - b) <= rtol * np.maximum(np.abs(a), np.abs(b))


def row_obeys_tie_rule(D0, I0, D1, I1, rtol=RTOL):
    """One row of the reference-rule result (D0, I0) against one row of ours (D1, I1).  -1 ids are unfilled slots."""
    f0, f1 = I0 >= 0, I1 >= 0
    if int(f0.sum()) != int(f1.sum()):
        return False
    d0, d1 = np.sort(D0[f0]), np.sort(D1[f1])
    if not _close(d0, d1, rtol).all():
        return False
    s0, s1 = dict(zip(I0[f0].tolist(), D0[f0].tolist())), dict(zip(I1[f1].tolist(), D1[f1].tolist()))
    only0 = sorted(v for k, v in s0.items() if k not in s1)
    only1 = sorted(v for k, v in s1.items() if k not in s0)
    if len(only0) != len(only1):
        return False
    if only0 and not _close(only0, only1, rtol).all():
        return False
    # an id present in both rows carries the same distance in both (same table, same code)
    common = [k for k in s0 if k in s1]
    if common and not _close([s0[k] for k in common], [s1[k] for k in common], rtol).all():
        return False
    return True


def tie_rule_report(D0, I0, D1, I1, rtol=RTOL):
    """Whole result sets.  Returns {rows, rows_identical, rows_differing, rows_violating, overlap, worst_rows}."""
    D0, I0, D1, I1 = (np.asarray(x) for x in (D0, I0, D1, I1))
    n, k = I0.shape
    same = (I0 == I1).all(axis=1)
    bad = []
    inter = 0
    for r in range(n):
        if same[r]:
            inter += int((I0[r] >= 0).sum()) if (I0[r] >= 0).any() else 0
            continue
        a, b = set(I0[r][I0[r] >= 0].tolist()), set(I1[r][I1[r] >= 0].tolist())
        inter += len(a & b)
        if not row_obeys_tie_rule(D0[r], I0[r], D1[r], I1[r], rtol):
            bad.append(r)
    filled = int((I0 >= 0).sum())
    return {"rows": int(n), "rows_identical": int(same.sum()), "rows_differing": int(n - same.sum()), "rows_violating": len(bad),
            "overlap": (inter / filled) if filled else 1.0, "worst_rows": bad[:8], "rtol": rtol}
