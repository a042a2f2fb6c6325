"""Parity protocol of SURVEY.md 8c for index lists: identical, or every differing (row, position) sits within the
reference's own fp32 rounding of a decision boundary when judged with the fp64 forward.

The detected set is decided by three comparisons (/root/reference/utils/mask2samples.py:9,16-21):
    nms        score == max of its 21-window                      -> boundary: |score - largest other score in the window|
    threshold  score >= th            (float th)                  -> boundary: |score - th|
    row max    score >= max of the suppressed row (th None / 0.0) -> boundary: |score - row max|
A fp32 forward (the reference's, or ours) carries rounding noise; a decision whose fp64 margin is smaller than that noise
may legitimately flip between two correct fp32 implementations.  The noise scale used here is the reference's own:
tol(row) = 4 * max |reference fp32 mask - fp64 mask| over that row.  Anything else is a bug and fails the test.

Test infrastructure only (uses oracle/torch_port.py, the CPU restatement of the reference).
"""
import numpy as np
import torch

from oracle import torch_port


def to_sd(npz_weights):
    return {k: torch.from_numpy(np.ascontiguousarray(v)) for k, v in npz_weights.items()}


def port_masks(sd, x_rows, double=False):
    """(n, L) float32 rows -> (n, 4L) reference-port masks on the CPU (fp32, or the fp64 re-run)."""
    x = torch.from_numpy(np.ascontiguousarray(x_rows)).unsqueeze(1)
    with torch.no_grad():
        if double:
            sd64 = {k: v.double() for k, v in sd.items()}
            return torch_port.stofnet_forward(sd64, x.double()).squeeze(1).numpy()
        return torch_port.stofnet_forward(sd, x).squeeze(1).numpy()


def index_set(idx):
    return set(map(tuple, np.asarray(idx).reshape(-1, 2).tolist()))


def classify_mismatches(sd, x_rows, got_idx, ref_idx, th, ref_masks32=None, window=20):
    """Symmetric difference of two (K, 2) [row, position] lists, each entry judged with the fp64 forward.

    Returns (n_diff, n_marginal, report): the test passes iff n_diff == n_marginal."""
    diff = sorted(index_set(got_idx) ^ index_set(ref_idx))
    if not diff:
        return 0, 0, []
    h = (window // 2 * 2 + 1) // 2
    rows = sorted({r for r, _ in diff})
    m64 = port_masks(sd, x_rows[rows], double=True)
    m32 = ref_masks32[rows] if ref_masks32 is not None else port_masks(sd, x_rows[rows])
    report, marginal = [], 0
    for r, p in diff:
        k = rows.index(r)
        row64 = m64[k]
        tol = 4.0 * float(np.abs(m32[k].astype(np.float64) - row64).max())
        tol = max(tol, 8 * np.finfo(np.float32).eps * float(np.abs(row64).max()))
        lo, hi = max(0, p - h), min(row64.shape[0], p + h + 1)
        others = np.delete(row64[lo:hi], p - lo)
        margins = {"nms": abs(row64[p] - others.max()) if others.size else np.inf}
        if th:
            margins["threshold"] = abs(row64[p] - th)
        else:
            # row max of the suppressed scores: the largest window maximum (or 0 when something is suppressed)
            margins["rowmax"] = abs(row64[p] - max(row64.max(), 0.0))
        which = min(margins, key=margins.get)
        ok = margins[which] <= tol
        marginal += int(ok)
        report.append({"row": int(r), "pos": int(p), "in": "ours" if (r, p) in index_set(got_idx) else "reference",
                       "score64": float(row64[p]), "boundary": which, "margin": float(margins[which]), "tol": tol,
                       "marginal": bool(ok)})
    return len(diff), marginal, report


def assert_index_parity(sd, x_rows, got_idx, ref_idx, th, ref_masks32=None, what=""):
    n, m, rep = classify_mismatches(sd, x_rows, got_idx, ref_idx, th, ref_masks32)
    if n:
        print("parity %s: %d differing (row, position) entries, %d within fp32 rounding of a decision boundary" % (what, n, m))
        for r in rep:
            print("   ", r)
    assert n == m, "%s: %d index mismatches that are NOT explained by a decision-boundary margin: %s" % (
        what, n - m, [r for r in rep if not r["marginal"]])
    return n
