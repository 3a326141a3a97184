"""SPRITE: the A-step (igm/steps/SpriteAssignmentStep.py) and the restraint (igm/restraints/sprite.py) for a whole
population.

    clusters (haploid segment ids)                                         restraints/sprite/clusters (indptr, data)
      -> ClusterBatch: segments grouped by chromosome, one representative per group     sprite.pyx:186-226
      -> Population.sprite_select: Rg^2 of every structure, the keep_best smallest      sprite.pyx:104-283, task :140-160
      -> assign_clusters: Gibbs draw of ONE structure per cluster with an occupancy penalty         reduce :176-272
      -> centroid_restraints: a DUMMY_DYNAMIC centroid per assigned cluster + bead-centroid bonds   sprite.py:41-69

The reference draws its random numbers from the global numpy state (np.random.choice for the representatives,
np.random.permutation / np.random.rand in the Gibbs step); here the generator is an argument (a
``numpy.random.RandomState``), used with the same calls in the same order, so a run can be replayed.
"""
import ctypes as C

import numpy as np

from . import _abi
from ._abi import KIND_CENTROID, BOND_UPPER
from .astep import Population

MAX_GROUPS = 8


class ClusterBatch(object):
    """Flat description of a batch of clusters for ``igmk_pop_sprite_select``.

    clusters: list of integer arrays (haploid segment ids); chrom_hap: chromosome of every haploid id; copy_off /
    copy_beads: Index.copy_index flattened.  Clusters on more than ``max_chrom`` chromosomes are skipped like
    SpriteAssignmentStep.task does (:124-131: filled with -1)."""

    def __init__(self, clusters, chrom_hap, copy_off, copy_beads, rng, max_chrom=6):
        chrom_hap = np.asarray(chrom_hap)
        self.n_clusters = len(clusters)
        self.skipped = np.zeros(self.n_clusters, dtype=bool)
        self.sizes = np.array([len(c) for c in clusters], dtype=np.int64)
        cl_off, seg_off, beads, seg_group, rep, self.kept = [0], [0], [], [], [], []
        self.segments = []
        for ci, cluster in enumerate(clusters):
            cluster = np.sort(np.asarray(cluster, dtype=np.int64))            # sprite.pyx:192
            cchroms = chrom_hap[cluster]
            uniq = np.unique(cchroms)
            if len(uniq) > max_chrom or len(uniq) > MAX_GROUPS:
                self.skipped[ci] = True
                continue
            if np.any(np.diff(cchroms) < 0):
                raise ValueError("haploid ids must be ordered by chromosome (alabtools Index)")
            groups = np.searchsorted(uniq, cchroms)
            r = np.full(MAX_GROUPS, -1, dtype=np.int32)
            if len(uniq) > 1:                                               # :218-223: np.random.choice per chromosome, in order
                for g in range(len(uniq)):
                    members = np.nonzero(groups == g)[0]
                    r[g] = members[int(np.nonzero(cluster[members] == rng.choice(cluster[members]))[0][0])]
            else:
                r[0] = 0
            for h, g in zip(cluster, groups):
                b = copy_beads[copy_off[h]:copy_off[h + 1]]
                beads.extend(int(v) for v in b)
                seg_off.append(len(beads))
                seg_group.append(int(g))
            cl_off.append(len(seg_off) - 1)
            rep.append(r)
            self.kept.append(ci)
            self.segments.append(cluster)
        self.cl_off = np.array(cl_off, dtype=np.int32)
        self.seg_off = np.array(seg_off, dtype=np.int32)
        self.beads = np.array(beads if beads else [0], dtype=np.int32)
        self.seg_group = np.array(seg_group if seg_group else [0], dtype=np.int32)
        self.rep = np.array(rep, dtype=np.int32).reshape(-1, MAX_GROUPS)
        self.kept = np.array(self.kept, dtype=np.int64)


def _sprite_select(self, batch, keep_best=50):
    """-> (best_rg2 f4 [n_clusters][keep], best_struct i4 [n_clusters][keep], selected: list per cluster of i4
    [keep][n_segments]) -- what SpriteAssignmentStep.task stores per cluster (values, indexes, selected beads); skipped
    clusters hold -1 everywhere (:124-131)."""
    nk = len(batch.kept)
    keep = int(keep_best)
    rg2 = np.full((batch.n_clusters, keep), -1.0, dtype=np.float32)
    st = np.full((batch.n_clusters, keep), -1, dtype=np.int32)
    selected = [np.full((keep, int(batch.sizes[c])), -1, dtype=np.int32) for c in range(batch.n_clusters)]
    if nk:
        v = np.empty((nk, keep), dtype=np.float32)
        s = np.empty((nk, keep), dtype=np.int32)
        bb = np.empty((int(batch.cl_off[-1]), keep), dtype=np.int32)
        _abi.check(self._lib.igmk_pop_sprite_select(
            self._p, nk, batch.cl_off.ctypes.data_as(_abi.c_i32p), batch.seg_off.ctypes.data_as(_abi.c_i32p),
            batch.beads.ctypes.data_as(_abi.c_i32p), batch.seg_group.ctypes.data_as(_abi.c_i32p),
            batch.rep.ctypes.data_as(_abi.c_i32p), keep, v.ctypes.data_as(_abi.c_f32p), s.ctypes.data_as(_abi.c_i32p),
            bb.ctypes.data_as(_abi.c_i32p)), "igmk_pop_sprite_select")
        rg2[batch.kept] = v
        st[batch.kept] = s
        for k, c in enumerate(batch.kept):
            selected[c] = np.ascontiguousarray(bb[batch.cl_off[k]:batch.cl_off[k + 1]].T)
    return rg2, st, selected


Population.sprite_select = _sprite_select


def assign_clusters(best_rg2, best_struct, n_struct, rng, batch_size=10, kT=100.0):
    """SpriteAssignmentStep.reduce (:176-272): the batches in a random order, every cluster draws one of its keep_best
    structures from a Gibbs distribution over the radius of gyration with a penalty for structures that already hold more
    than the average number of clusters.  -> (assignment i4 [n_clusters] (-1: skipped), pos i4 [n_clusters]: which of
    the keep_best entries was drawn)"""
    best_rg2 = np.asarray(best_rg2)
    n_clusters = best_rg2.shape[0]
    n_batches = n_clusters // batch_size + (1 if n_clusters % batch_size else 0)
    order = rng.permutation(range(n_batches))                                   # :179
    occupancy = np.zeros(n_struct, dtype=np.int32)
    assignment = np.zeros(n_clusters, dtype=np.int32)
    pos_out = np.zeros(n_clusters, dtype=np.int32)
    aveN = float(n_clusters) / n_struct
    stdN = np.sqrt(aveN)
    for batch_id in order:
        for ci in range(batch_id * batch_size, min(n_clusters, (batch_id + 1) * batch_size)):
            rg2s, curr_idx = best_rg2[ci], best_struct[ci]
            if rg2s[0] < 0:                                                      # :233-235
                pos, si = 0, -1
            else:
                best_rgs = np.sqrt(rg2s)
                pen = np.clip(occupancy[curr_idx] - aveN, 0., None) / stdN
                E = (best_rgs - best_rgs[0]) / kT + pen
                P = np.cumsum(np.exp(-(E - E[0])))
                e = rng.rand() * P[-1]
                pos = int(np.searchsorted(P, e, side='left'))
                si = int(curr_idx[pos])
                occupancy[si] += 1
            assignment[ci] = si
            pos_out[ci] = pos
    return assignment, pos_out


def get_cluster_size(radii, volume_occupancy):
    """sprite.py:72-76"""
    return (np.sum(radii ** 3) / volume_occupancy) ** (1. / 3.)


def centroid_restraints(crd, radii, assignment, selected_beads, s_begin, first_atom, volume_occupancy, k=1.0):
    """Sprite._apply (sprite.py:41-69) for structures s_begin .. s_begin + S_local - 1 of a population.

    crd: float32 [n_bead][S_local][3]; assignment [n_clusters]: global structure index of every cluster;
    selected_beads: list per cluster of the bead ids chosen for its assigned structure; first_atom: index the first
    centroid of a structure gets (= number of atoms of the template).
    -> (atoms (off i8 [S+1], xyz f4, radius f4, kind u1), bonds (off, i, j, k, r0, style)) for PopulationExtras."""
    crd = np.asarray(crd, dtype=np.float32)
    S = crd.shape[1]
    radii = np.asarray(radii, dtype=np.float32)
    a_off, b_off = [0], [0]
    xyz, bi, bj, br0 = [], [], [], []
    assignment = np.asarray(assignment)
    for s in range(S):
        n_c = 0
        for ci in np.where(assignment == s_begin + s)[0]:                         # :49
            beads = np.asarray(selected_beads[ci], dtype=np.int64)
            crad = radii[beads]
            csize = get_cluster_size(crad, volume_occupancy)                      # :56
            xyz.append(np.mean(crd[beads, s, :], axis=0))                         # :59 (float32 mean -> float32 position)
            cen = first_atom + n_c
            n_c += 1
            for b in beads:                                                       # :62-67
                bi.append(int(b)); bj.append(cen); br0.append(float(csize - radii[b]))
        a_off.append(a_off[-1] + n_c)
        b_off.append(len(bi))
    na, nb = a_off[-1], len(bi)
    atoms = (np.array(a_off, dtype=np.int64), np.array(xyz, dtype=np.float32).reshape(na, 3), np.zeros(na, dtype=np.float32),
             np.full(na, KIND_CENTROID, dtype=np.uint8))
    bonds = (np.array(b_off, dtype=np.int64), np.array(bi, dtype=np.int32), np.array(bj, dtype=np.int32),
             np.full(nb, float(k), dtype=np.float64), np.array(br0, dtype=np.float64), np.full(nb, BOND_UPPER, dtype=np.uint8))
    return atoms, bonds
