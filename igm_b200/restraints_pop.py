"""Population-wide builders of the restraints that differ from structure to structure beyond the Hi-C rows:
what ``Fish._apply`` (igm/restraints/fish.py:93-274) and ``Damid._apply_envelope`` (igm/restraints/damid.py:116-142)
add to ONE Model inside ``ModelingStep.task`` (igm/steps/ModelingStep.py:268-330), here for all structures of a
population at once, as flat arrays for ``igmk_submit_population_ex`` (no Python objects per force).

Host logic only (numpy): the selections depend on one structure's own coordinates and cost O(probes + pairs) each.
"""
import numpy as np

from ._abi import BOND_UPPER, BOND_LOWER
from .kernel import PopulationExtras


def particle_norm(v):
    """``numpy.linalg.norm`` of float32 3-vectors the way the reference evaluates it for ONE vector
    (fish.py:30,44 ``norm(crd[i])``, ``norm(x - y)``): sqrt(dot(v, v)) through BLAS sdot -- float32 products, a double
    accumulator in index order, one rounding to float32, float32 sqrt.  Vectorised over leading axes."""
    v = np.asarray(v, dtype=np.float32)
    sq = (v * v).astype(np.float32)
    acc = (sq[..., 0].astype(np.float64) + sq[..., 1].astype(np.float64)) + sq[..., 2].astype(np.float64)
    return np.sqrt(acc.astype(np.float32))


def _copies(copy_off, copy_beads, h):
    return np.asarray(copy_beads[copy_off[h]:copy_off[h + 1]], dtype=np.int64)


def fish_bonds(crd, copy_off, copy_beads, center, rtype, tol, k, probes=None, radial_min=None, radial_max=None,
               pairs=None, pair_min=None, pair_max=None):
    """Fish._apply for every structure of ``crd`` (float32 [n_bead][S][3]).

    probes: haploid indices [nq]; radial_min / radial_max: float64 [nq][S], the targets the FISH assignment gave each
    structure (``hff['radial_min'][:, struct_id]``, fish.py:124); pairs: [np][2]; pair_min / pair_max: [np][S].
    center: atom index of the restraint's DUMMY_STATIC particle (fish.py:119).  rtype: subset of "rRpP".
    -> (off i8 [S+1], i i4, j i4, k f8, r0 f8, style u1): per-structure bonds in the order the reference adds them."""
    crd = np.asarray(crd, dtype=np.float32)
    S = crd.shape[1]
    bi, bj, br0, bs = [], [], [], []          # lists of [S] columns, one per bond of a structure

    def add(i, j, d, style):
        bi.append(np.broadcast_to(np.asarray(i, dtype=np.int64), (S,)))
        bj.append(np.broadcast_to(np.asarray(j, dtype=np.int64), (S,)))
        br0.append(np.broadcast_to(np.asarray(d, dtype=np.float64), (S,)))
        bs.append(style)

    def radial_order(h):
        ii = _copies(copy_off, copy_beads, h)
        d = particle_norm(crd[ii])                                   # [copies][S]
        order = np.argsort(d, axis=0, kind="stable")                 # sort_radially, fish.py:26-32
        return ii[order[0]], ii[order[-1]]                           # closest to / farthest from the centre, per structure

    def pair_order(hi, hj):
        ii, jj = _copies(copy_off, copy_beads, hi), _copies(copy_off, copy_beads, hj)
        m = np.repeat(ii, jj.size)
        n = np.tile(jj, ii.size)                                     # for m in ii: for n in jj (fish.py:41-47)
        d = particle_norm(crd[m] - crd[n]).astype(np.float64)        # [comb][S]
        order = np.argsort(d, axis=0, kind="stable")
        return m[order], n[order]                                    # [comb][S], sorted by distance per structure

    if "r" in rtype or "R" in rtype:
        if probes is None:
            raise ValueError("radial FISH restraints need probes")
    if "r" in rtype:                                                 # fish.py:121-159
        for q, h in enumerate(probes):
            t = np.asarray(radial_min[q], dtype=np.float64)
            near, far = radial_order(int(h))
            add(center, near, np.maximum(0, t - tol), BOND_LOWER)
            add(center, near, t + tol, BOND_UPPER)
            add(center, far, t + tol, BOND_LOWER)
    if "R" in rtype:                                                 # fish.py:162-196
        for q, h in enumerate(probes):
            t = np.asarray(radial_max[q], dtype=np.float64)
            near, far = radial_order(int(h))
            add(center, far, np.maximum(0, t - tol), BOND_LOWER)
            add(center, far, t + tol, BOND_UPPER)
            add(center, near, t - tol, BOND_UPPER)
    if "p" in rtype:                                                 # fish.py:198-235
        for q, (hi, hj) in enumerate(pairs):
            if hi == hj:
                raise AssertionError("FISH pair with i == j")        # fish.py:204
            t = np.asarray(pair_min[q], dtype=np.float64)
            m, n = pair_order(int(hi), int(hj))
            for c in range(m.shape[0]):
                add(m[c], n[c], np.maximum(0, t - tol), BOND_LOWER)
            add(m[0], n[0], t + tol, BOND_UPPER)
    if "P" in rtype:                                                 # fish.py:237-272
        for q, (hi, hj) in enumerate(pairs):
            if hi == hj:
                raise AssertionError("FISH pair with i == j")
            t = np.asarray(pair_max[q], dtype=np.float64)
            m, n = pair_order(int(hi), int(hj))
            for c in range(m.shape[0]):
                add(m[c], n[c], t + tol, BOND_UPPER)
            add(m[-1], n[-1], np.maximum(0, t - tol), BOND_LOWER)
    nb = len(bi)
    off = (np.arange(S + 1, dtype=np.int64) * nb)
    if nb == 0:
        z = np.zeros(0)
        return off, z.astype(np.int32), z.astype(np.int32), z, z, z.astype(np.uint8)
    out_i = np.stack(bi, axis=1).astype(np.int32).ravel()            # [S][nb] -> structure-major
    out_j = np.stack(bj, axis=1).astype(np.int32).ravel()
    out_r0 = np.stack(br0, axis=1).ravel()
    out_s = np.tile(np.array(bs, dtype=np.uint8), S)
    return off, out_i, out_j, np.full(out_i.size, float(k), dtype=np.float64), out_r0, out_s


def fish_targets(ranks, sorted_targets):
    """What FishAssignmentStep.task stores per item (FishAssignmentStep.py:220-225, :247-252): ``target[idx]`` -- the
    experimental distance distribution of the item (one value per structure of the WHOLE population, as given in the
    input file) permuted by the rank of every structure.  ranks: i4 [n][S]; sorted_targets: [n][S].  -> f8 [n][S]"""
    t = np.asarray(sorted_targets, dtype=np.float64)
    r = np.asarray(ranks, dtype=np.int64)
    if t.shape != r.shape:
        raise ValueError("one target value per structure is needed (FishAssignmentStep.py:220)")
    return np.take_along_axis(t, r, axis=1)


def damid_envelope(bead, bits, semiaxes, contact_range, k):
    """The group Damid._apply_envelope adds (damid.py:131-139): EllipticEnvelope(affected particles of this structure,
    semiaxes * (1 - contact_range), -k).  bead / bits: rows of the DamID actdist file and the assignment matrix
    Population.damid_assign returned for exactly the structures being submitted.  -> an `envelopes` entry of PopulationExtras"""
    abc = np.asarray(semiaxes, dtype=np.float64) * (1 - contact_range)
    return ((float(abc[0]), float(abc[1]), float(abc[2]), -float(k)), np.asarray(bead, dtype=np.int32), np.asarray(bits, dtype=np.uint32))


def extras(n_struct, bonds=None, envelopes=(), atoms=None):
    return PopulationExtras(n_struct, bonds=bonds, envelopes=envelopes, atoms=atoms)
