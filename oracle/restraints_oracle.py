"""CPU restatement of what ModelingStep.task puts on ONE Model for the DamID and FISH restraints.
TEST INFRASTRUCTURE ONLY (tests/, never the product path).

Follows igm/restraints/fish.py:93-274 (Fish._apply), :26-50 (sort_radially, sort_pairs_by_distance) and
igm/restraints/damid.py:116-142 (Damid._apply_envelope) statement by statement, with plain Python loops.
Pinned by tests/golden/restraints_golden.npz: the force lists the reference's own classes built on a Model
(tests/golden/make_golden_restraints.py)."""
import numpy as np

from .astep_oracle import particle_distance
from . import damid_oracle

UPPER, LOWER = 0, 1


def _norm(v):
    return particle_distance(np.asarray(v, dtype=np.float32), np.zeros(3, dtype=np.float32))


def sort_radially(ii, crd):
    ii = np.array(ii)
    dists = [_norm(crd[i]) for i in ii]
    return ii[np.argsort(dists)]


def sort_pairs_by_distance(ii, jj, crd):
    n_combinations = len(ii) * len(jj)
    dists = np.zeros(n_combinations)
    pairs = np.zeros((n_combinations, 2), dtype=int)
    it = 0
    for m in ii:
        for n in jj:
            dists[it] = particle_distance(crd[m], crd[n])
            pairs[it, :] = [m, n]
            it += 1
    return pairs[np.argsort(dists), :]


def fish_forces(crd, copy_index, center, rtype, tol, ck, struct_id, probes=None, radial_min=None, radial_max=None,
                pairs=None, pair_min=None, pair_max=None):
    """crd: (n_bead, 3) float32 = model.getCoordinates() of THIS structure.  -> [(style, i, j, d, k)] in the order
    Fish._apply adds the forces."""
    out = []
    if "r" in rtype:
        targets = np.asarray(radial_min)[:, struct_id]
        for q, i in enumerate(probes):
            ii = copy_index[int(i)]
            target_dist = targets[q]
            particle = sort_radially(ii, crd)[0]
            out.append((LOWER, center, particle, max(0, target_dist - tol), ck))
            out.append((UPPER, center, particle, target_dist + tol, ck))
            other_particle = sort_radially(ii, crd)[-1]
            out.append((LOWER, center, other_particle, target_dist + tol, ck))
    if "R" in rtype:
        targets = np.asarray(radial_max)[:, struct_id]
        for q, i in enumerate(probes):
            ii = copy_index[int(i)]
            target_dist = targets[q]
            particle = sort_radially(ii, crd)[-1]
            out.append((LOWER, center, particle, max(0, target_dist - tol), ck))
            out.append((UPPER, center, particle, target_dist + tol, ck))
            other_particle = sort_radially(ii, crd)[0]
            out.append((UPPER, center, other_particle, target_dist - tol, ck))
    if "p" in rtype:
        targets = np.asarray(pair_min)[:, struct_id]
        for q, (i, j) in enumerate(pairs):
            assert i != j
            target_dist = targets[q]
            sorted_pairs = sort_pairs_by_distance(copy_index[int(i)], copy_index[int(j)], crd)
            for m, n in sorted_pairs:
                out.append((LOWER, m, n, max(0, target_dist - tol), ck))
            m, n = sorted_pairs[0]
            out.append((UPPER, m, n, target_dist + tol, ck))
    if "P" in rtype:
        targets = np.asarray(pair_max)[:, struct_id]
        for q, (i, j) in enumerate(pairs):
            assert i != j
            target_dist = targets[q]
            sorted_pairs = sort_pairs_by_distance(copy_index[int(i)], copy_index[int(j)], crd)
            for m, n in sorted_pairs:
                out.append((UPPER, m, n, target_dist + tol, ck))
            m, n = sorted_pairs[-1]
            out.append((LOWER, m, n, max(0, target_dist - tol), ck))
    return out


def damid_members(crd_struct, radii, bead, dist, contact_range, semiaxes):
    """affected_particles of Damid._apply_envelope (damid.py:122-129) for one structure: (n_bead, 3) coordinates"""
    mask = damid_oracle.damid_assign(crd_struct[:, None, :], radii, bead, dist, contact_range, semiaxes)
    return [int(b) for b, m in zip(bead, mask[:, 0]) if m]
