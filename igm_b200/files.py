"""The population files either side of the hot path, without h5py / alabtools (SURVEY §8 f, rank 2):

    .hss           alabtools HssFile: `coordinates` (nbead, nstruct, 3) f4 -- the layout the kernels read --,
                   `radii`, `index/{chrom, copy, start, end, label, chrom_sizes, copy_index}`, `genome/*`,
                   `envelope/*`, JSON strings `summary`, `config_data`; root attributes nbead, nstruct,
                   violation, version                      (igm/steps/ModelingStep.py:441,517-527)
    .hcs           alabtools Contactmatrix: `matrix/{data f4, indices i4, indptr i4, diagonal f4}` = CSR of the
                   upper triangle over haploid bins, + `index`, `genome`
                                                            (igm/steps/ActivationDistanceStep.py:121-123)
    actdist.hdf5   `row i4, col i4, dist f4, prob f4`       (igm/steps/ActivationDistanceStep.py:277-281)

Reading goes through ``h5lite.File`` (pinned on the h5py-written demo files of the reference); writing through
``h5lite.write`` (contiguous, uncompressed; read back with the same reader -- libhdf5 is not available here to
open the result, see h5lite.py).
"""
import json

import numpy as np

from . import h5lite


class Index(object):
    """What the hot path reads from alabtools' Index: chrom, copy, copy_index (+ its CSR form)."""

    def __init__(self, chrom, copy, copy_index):
        self.chrom = np.ascontiguousarray(chrom, dtype=np.int32)
        self.copy = np.ascontiguousarray(copy, dtype=np.int32)
        self.copy_index = {int(k): [int(b) for b in v] for k, v in copy_index.items()}
        self.n_bead = int(self.chrom.size)
        self.n_hap = len(self.copy_index)
        off = np.zeros(self.n_hap + 1, dtype=np.int32)
        beads = []
        for h in range(self.n_hap):
            beads += self.copy_index[h]
            off[h + 1] = len(beads)
        self.copy_off = off
        self.copy_beads = np.array(beads, dtype=np.int32)

    def __len__(self):
        return self.n_bead


def _index_of(group):
    return Index(group["chrom"].read(), group["copy"].read(), json.loads(group["copy_index"].read()))


def read_hss(path):
    """-> dict(coordinates f4 [nbead][nstruct][3], radii f4, index Index, nbead, nstruct, tree): `tree` is the whole
    file as ``h5lite.read_tree`` returns it, for ``write_hss``."""
    tree = h5lite.read_tree(path)
    f = h5lite.File(path)
    crd = np.ascontiguousarray(tree["coordinates"], dtype=np.float32)
    attrs = tree.get(h5lite.ATTRS, {})
    nbead, nstruct = int(attrs.get("nbead", crd.shape[0])), int(attrs.get("nstruct", crd.shape[1]))
    if crd.shape != (nbead, nstruct, 3):
        raise ValueError("%s: coordinates %s do not match nbead=%d nstruct=%d" % (path, crd.shape, nbead, nstruct))
    return {"coordinates": crd, "radii": np.ascontiguousarray(tree["radii"], dtype=np.float32), "index": _index_of(f["index"]),
            "nbead": nbead, "nstruct": nstruct, "tree": tree}


def write_hss(path, tree, coordinates=None, summary=None, violation=None):
    """Write a population file: `tree` (from ``read_hss``) with, optionally, new coordinates / summary JSON /
    violation score -- what ModelingStep.reduce leaves behind (ModelingStep.py:517-547)."""
    out = dict(tree)
    attrs = dict(out.get(h5lite.ATTRS, {}))
    if coordinates is not None:
        crd = np.ascontiguousarray(coordinates, dtype=np.float32)
        if crd.ndim != 3 or crd.shape[2] != 3:
            raise ValueError("coordinates must be [nbead][nstruct][3]")
        out["coordinates"] = crd
        attrs["nbead"], attrs["nstruct"] = np.int64(crd.shape[0]), np.int64(crd.shape[1])
    if summary is not None:
        out["summary"] = summary if isinstance(summary, str) else json.dumps(summary)
    if violation is not None:
        attrs["violation"] = np.float64(violation)
    out[h5lite.ATTRS] = attrs
    h5lite.write(path, out)


def read_hcs(path):
    """-> dict(data f4, indices i4, indptr i4, diagonal f4, index Index, n): CSR upper triangle over haploid bins"""
    f = h5lite.File(path)
    m = f["matrix"]
    out = {k: m[k].read() for k in ("data", "indices", "indptr", "diagonal")}
    out["index"] = _index_of(f["index"])
    out["n"] = int(out["indptr"].size - 1)
    return out


def read_actdist(path):
    f = h5lite.File(path)
    return tuple(f[k].read() for k in ("row", "col", "dist", "prob"))


def write_actdist(path, row, col, dist, prob):
    """ActivationDistanceStep.reduce's output file (:277-281)"""
    h5lite.write(path, {"row": np.asarray(row, dtype=np.int32), "col": np.asarray(col, dtype=np.int32),
                        "dist": np.asarray(dist, dtype=np.float32), "prob": np.asarray(prob, dtype=np.float32)})


def hic_pairs_from_matrix(hcs, intra_sigma, inter_sigma, last_actdist=None):
    """The argument batches of ActivationDistanceStep.setup (:139-176), concatenated: every stored entry (i, j, pwish)
    of the matrix with pwish >= intra_sigma (same chromosome) or >= inter_sigma (different), in CSR order, with the
    corrected probability of the previous iteration: the `prob` of the last actdist file at (row, col) = (i, j) --
    rows whose beads are both first copies, duplicates summed as coo -> lil does (:141-153).  A sigma of False / None
    switches that class off.  The diagonal is not enumerated: i == j never yields a row (:351-352).
    -> (pair_i i4, pair_j i4, pwish f8, plast f8)."""
    n = hcs["n"]
    indptr, indices, data = hcs["indptr"], hcs["indices"], hcs["data"]
    i = np.repeat(np.arange(n, dtype=np.int32), np.diff(indptr))
    j = indices.astype(np.int32)
    chrom = hcs["index"].chrom
    same = chrom[i] == chrom[j]
    keep = np.zeros(i.size, dtype=bool)
    # `pwish >= sigma` with pwish a float32 scalar and sigma a Python float: float32 comparison under numpy >= 2
    # (NEP 50), the convention of the whole A-step here -- an entry equal to float32(sigma) is kept
    if intra_sigma is not None and intra_sigma is not False:
        keep |= same & (data >= np.float32(intra_sigma))
    if inter_sigma is not None and inter_sigma is not False:
        keep |= ~same & (data >= np.float32(inter_sigma))
    keep &= i != j
    pi, pj, pw = i[keep], j[keep], data[keep].astype(np.float64)
    pl = np.zeros(pi.size, dtype=np.float64)
    if last_actdist is not None:
        row, col, _, prob = last_actdist
        ok = (row < n) & (col < n)
        key = row[ok].astype(np.int64) * n + col[ok]
        uk, inv = np.unique(key, return_inverse=True)
        acc = np.zeros(uk.size, dtype=np.float64)
        np.add.at(acc, inv, prob[ok].astype(np.float64))
        want = pi.astype(np.int64) * n + pj
        pos = np.searchsorted(uk, want)
        pos[pos >= uk.size] = 0
        hit = uk.size > 0
        if hit:
            found = uk[pos] == want
            pl[found] = acc[pos[found]]
    return pi, pj, pw, pl
