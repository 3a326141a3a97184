"""CPU restatement (numpy) of the per-restraint violation ratios (M10).  TEST INFRASTRUCTURE ONLY.
Follows igm/model/forces.py:128-136, :176-184 (harmonic bounds) and :220-250 (EllipticEnvelope
.getViolationRatios); pinned by tests/golden/viol_golden.npz (outputs of the reference's classes)."""
import numpy as np

from .astep_oracle import particle_distance


def bond_ratios(xyz, bi, bj, r0, k, style):
    """xyz float32 [n][3]; r0/k float64 arrays; style 0 upper / 1 lower.  A bound distance that is a
    float32 value takes numpy's float32 chain, any other the float64 chain (see mstep.cuh)."""
    out = np.zeros(len(bi))
    dist = particle_distance(xyz[bi], xyz[bj])                      # float32
    for n in range(len(bi)):
        d, kk = r0[n], k[n]
        if d == 0:
            continue
        if float(np.float32(d)) == d:
            d, kk, x = np.float32(d), np.float32(kk), dist[n]
        else:
            d, kk, x = np.float64(d), np.float64(kk), np.float64(dist[n])
        if style[n] == 0:
            score = 0 if x <= d else kk * (x - d)
        else:
            score = 0 if x >= d else kk * (d - x)
        out[n] = score / (kk * d)
    return out


def envelope_u(xyz, radius, members, semiaxes, k):
    """scale-free part of EllipticEnvelope.getScores (forces.py:224-244)"""
    out = np.zeros(len(members))
    for n, i in enumerate(members):
        s2 = np.square(np.asarray(semiaxes, dtype=np.float64) - np.float32(radius[i]))
        x = xyz[i]
        q = np.sqrt(np.sum(x ** 2 / s2))
        if q > 1 and k > 0:
            out[n] = (1.0 - 1.0 / np.sqrt(q)) * np.linalg.norm(x)
        elif q < 1 and k < 0:
            out[n] = 1.0 - np.sqrt(q)
    return out
