"""CPU restatement (numpy) of the FISH A-step helpers.  TEST INFRASTRUCTURE ONLY.
Follows igm/steps/FishAssignmentStep.py:31-88; pinned by tests/golden/fish_golden.npz (outputs of
the reference's own get_pair_dists / get_rad_dists / get_min_max_and_idx)."""
import numpy as np


def _norm(d):
    """np.linalg.norm(axis=1) on float32 rows = sqrt((x2+y2)+z2) in float32"""
    sq = (d * d).astype(np.float32)
    return np.sqrt((sq[:, 0] + sq[:, 1]) + sq[:, 2])


def fish_ranks(crd, copy_index, hap_i, hap_j=None):
    """-> mindists, maxdists (float32 values), idxmin, idxmax per item (ties ranked by structure index)"""
    lo, hi, ra, rb = [], [], [], []
    for q, i in enumerate(hap_i):
        ii = copy_index[int(i)]
        if hap_j is not None and hap_j[q] >= 0:
            jj = copy_index[int(hap_j[q])]
            d = np.stack([_norm(crd[ii[0]] - crd[jj[0]]), _norm(crd[ii[0]] - crd[jj[1]]),
                          _norm(crd[ii[1]] - crd[jj[0]]), _norm(crd[ii[1]] - crd[jj[1]])])
        else:
            d = np.stack([_norm(crd[ii[0]]), _norm(crd[ii[1]])])
        mn, mx = d.min(axis=0), d.max(axis=0)
        lo.append(mn); hi.append(mx)
        ra.append(np.argsort(np.argsort(mn, kind="stable"), kind="stable"))
        rb.append(np.argsort(np.argsort(mx, kind="stable"), kind="stable"))
    return np.array(lo), np.array(hi), np.array(ra, dtype=np.int32), np.array(rb, dtype=np.int32)
