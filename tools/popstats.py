#!/usr/bin/env python
"""Structural statistics of population files (.hss), to compare a population made here with one the reference
recorded: contact probability and mean distance by genomic separation, radial position, territory size, and --
with a contact matrix -- the contact frequency of the population next to the Hi-C target per separation class.

    python tools/popstats.py [--hcs matrix.hcs --sigma 0.02] a.hss [b.hss ...]"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from igm_b200 import files          # noqa: E402


def structure_stats(pop):
    crd = pop["coordinates"].astype(np.float64); idx = pop["index"]; r = float(pop["radii"][0])
    n = idx.n_bead
    out = []
    for s in (1, 2, 3, 5, 10, 20):
        ok = (idx.chrom[:-s] == idx.chrom[s:]) & (idx.copy[:-s] == idx.copy[s:])
        d = np.linalg.norm(crd[:-s][ok] - crd[s:][ok], axis=2)
        out.append(("P(contact), separation %d" % s, float((d <= 4 * r).mean())))
        out.append(("mean distance / r, separation %d" % s, float(d.mean() / r)))
    R = float(np.asarray(pop["tree"]["envelope"]["params"]).ravel()[0])
    out.append(("mean radial position / R", float(np.linalg.norm(crd, axis=2).mean() / R)))
    rng = np.random.default_rng(0)
    a, b, s_ = rng.integers(0, n, 200000), rng.integers(0, n, 200000), rng.integers(0, crd.shape[1], 200000)
    d = np.linalg.norm(crd[a, s_] - crd[b, s_], axis=1)[a != b]
    out.append(("random bead pairs: P(d < 2r)", float((d < 2 * r).mean())))
    out.append(("random bead pairs: P(d < 4r)", float((d < 4 * r).mean())))
    m = (idx.chrom == 0) & (idx.copy == 0)
    c = crd[m]
    out.append(("radius of gyration of chr1 copy 0 / r", float(np.sqrt(((c - c.mean(axis=0)) ** 2).sum(axis=2).mean(axis=0)).mean() / r)))
    return out


def contact_vs_target(pop, hcs, sigma, sample=3000):
    """pnow of get_actdist (ActivationDistanceStep.py:419-422) for a sample of the pairs kept at sigma, by class"""
    crd, idx, r = pop["coordinates"], pop["index"], pop["radii"]
    pi, pj, pw, _ = files.hic_pairs_from_matrix(hcs, sigma, sigma)
    sel = np.random.default_rng(0).choice(pi.size, min(sample, pi.size), replace=False)
    pi, pj, pw = pi[sel], pj[sel], pw[sel]
    pn = np.zeros(pi.size)
    for q, (i, j) in enumerate(zip(pi, pj)):
        ii, jj = idx.copy_index[int(i)], idx.copy_index[int(j)]
        pairs = list(zip(ii, jj)) if idx.chrom[i] == idx.chrom[j] else [(a, b) for a in ii for b in jj]
        rc = 2.0 * (r[ii[0]] + r[jj[0]])
        pn[q] = sum((np.linalg.norm(crd[a] - crd[b], axis=1) <= rc).sum() for a, b in pairs) / (len(pairs) * crd.shape[1])
    ch = hcs["index"].chrom
    same, sep = ch[pi] == ch[pj], pj - pi
    out = []
    for lo, hi in ((1, 1), (2, 2), (3, 3), (4, 6), (7, 15), (16, 60), (61, 100000)):
        k = same & (sep >= lo) & (sep <= hi)
        if k.sum():
            out.append(("intra, separation %d-%d" % (lo, hi), int(k.sum()), float(pw[k].mean()), float(pn[k].mean())))
    k = ~same
    if k.sum():
        out.append(("inter", int(k.sum()), float(pw[k].mean()), float(pn[k].mean())))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("hss", nargs="+"); ap.add_argument("--hcs", default=None); ap.add_argument("--sigma", type=float, default=0.02)
    a = ap.parse_args()
    hcs = files.read_hcs(a.hcs) if a.hcs else None
    for path in a.hss:
        pop = files.read_hss(path)
        print(path)
        for k, v in structure_stats(pop):
            print("   %-42s %.4f" % (k, v))
        if hcs is not None:
            print("   contact frequency of the pairs kept at sigma %g (sample):   n   mean pwish   mean pnow" % a.sigma)
            for name, n, pw, pn in contact_vs_target(pop, hcs, a.sigma):
                print("   %-42s %5d   %.3f        %.3f" % (name, n, pw, pn))


if __name__ == "__main__":
    main()
