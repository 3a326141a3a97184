"""CPU restatement (numpy) of IGM's DamID (lamina) A-step.  TEST INFRASTRUCTURE ONLY.

Follows igm/steps/DamidActivationDistanceStep.py:35-76 (snormsq_*), :285-356
(get_damid_actdist), :31/:229 (text round trip) and igm/restraints/damid.py:55-75,116-130 (the
assignment predicate).  Pinned by tests/golden/damid_golden.npz, produced by the reference's own
functions (tests/golden/make_golden.py).  numpy >= 2 promotion rules are part of the definition:
float32 squares, float64 division by a numpy-scalar ``(axis - r)**2`` (scalar pow)."""
import numpy as np

from .astep_oracle import clean_probability, text_roundtrip


def _den(axis, r):
    return (np.float64(axis) - np.float32(r)) ** 2           # scalar `**` = libm pow


def snormsq_sphere(x, R, r):
    return np.sum(np.square(x), axis=1) / _den(R, r)


def snormsq_ellipse(x, semiaxes, r):
    sq = np.square(x)
    return sq[:, 0] / _den(semiaxes[0], r) + sq[:, 1] / _den(semiaxes[1], r) + sq[:, 2] / _den(semiaxes[2], r)


def get_damid_actdist(locid, pwish, plast, crd, radii, copy_index, contact_range=0.05, shape="sphere",
                      nucleus_param=5000.0):
    """-> [(bead, ad, p)] like DamidActivationDistanceStep.py:285-356 (before the text round trip)"""
    ii = copy_index[locid]
    n_struct = crd.shape[1]
    r = radii[ii[0]]
    R = np.array(nucleus_param) * (1 - contact_range)
    d_sq = np.empty(len(ii) * n_struct)
    for k, b in enumerate(ii):
        d_sq[k * n_struct:(k + 1) * n_struct] = snormsq_sphere(crd[b], R, r) if shape == "sphere" else snormsq_ellipse(crd[b], R, r)
    d_sq[::-1].sort()
    pnow = float(np.count_nonzero(d_sq >= 1.0)) / n_struct / len(ii)
    t = clean_probability(pnow, plast)
    p = clean_probability(pwish, t)
    ad = 2
    if p > 0:
        o = min(len(ii) * n_struct - 1, int(round(len(ii) * n_struct * p)))
        ad = np.sqrt(d_sq[o])
    return [(b, ad, p) for b in ii]


def damid_rows(loc, pwish, plast, crd, radii, copy_index, contact_range, shape, nucleus_param):
    bead, dist, prob = [], [], []
    for i, pw, pl in zip(loc, pwish, plast):
        for (b, ad, p) in get_damid_actdist(int(i), float(pw), float(pl), crd, radii, copy_index, contact_range, shape, nucleus_param):
            bead.append(b)
            dist.append(text_roundtrip(ad, 2))
            prob.append(text_roundtrip(p, 5))
    return np.array(bead, dtype=np.int32), np.array(dist, dtype=np.float32), np.array(prob, dtype=np.float32)


def damid_assign(crd, radii, bead, dist, contact_range, semiaxes):
    """Damid._apply_envelope's predicate (damid.py:122-129) for every (row, structure) -> bool [rows, S]"""
    abc = np.array(semiaxes, dtype=np.float64) * (1 - contact_range)
    out = np.zeros((len(bead), crd.shape[1]), dtype=bool)
    for k, (b, d) in enumerate(zip(bead, dist)):
        out[k] = snormsq_ellipse(crd[b], abc, radii[b]) >= np.float32(d) ** 2
    return out
