"""A-step host side: a device-resident population and the two Hi-C A-step operations.

``Population.actdist`` is ``ActivationDistanceStep.task`` + the parse in ``reduce``
(igm/steps/ActivationDistanceStep.py:192-222, :238-249) for one batch of haploid pairs;
``Population.assign`` evaluates the contact-assignment predicate of
``interHiC._apply`` / ``intraHiC._apply`` (igm/restraints/inter_hic.py:49-64,
intra_hic.py:50-65) for every (actdist row, structure).  Both run on the GPU through the C ABI
and fail loudly without it.
"""
import ctypes as C

import numpy as np

from . import _abi


def copy_index_to_csr(copy_index, n_hap=None):
    """Index.copy_index (dict haploid bin -> list of bead ids) -> (copy_off, copy_beads)."""
    if n_hap is None:
        n_hap = len(copy_index)
    off = np.zeros(n_hap + 1, dtype=np.int32)
    beads = []
    for h in range(n_hap):
        ids = copy_index[h] if h in copy_index else copy_index[str(h)]
        beads.extend(int(b) for b in ids)
        off[h + 1] = len(beads)
    return off, np.array(beads, dtype=np.int32)


class Population(object):
    """coordinates[bead][struct][3] float32 (the .hss layout) resident on one GPU."""

    def __init__(self, n_struct, radii, chrom, copy_off, copy_beads, device=0):
        self._lib = _abi.load()
        self.radii = np.ascontiguousarray(radii, dtype=np.float32)
        self.chrom = np.ascontiguousarray(chrom, dtype=np.int32)
        self.copy_off = np.ascontiguousarray(copy_off, dtype=np.int32)
        self.copy_beads = np.ascontiguousarray(copy_beads, dtype=np.int32)
        self.n_bead = int(self.radii.size)
        self.n_struct = int(n_struct)
        self.device = int(device)
        self.n_hap = int(self.copy_off.size - 1)
        self.max_copies = int(np.diff(self.copy_off).max())
        self._keep = None
        self._p = C.c_void_p()
        _abi.check(self._lib.igmk_pop_create(
            int(device), self.n_bead, self.n_struct, self.n_hap,
            self.radii.ctypes.data_as(_abi.c_f32p), self.chrom.ctypes.data_as(_abi.c_i32p),
            self.copy_off.ctypes.data_as(_abi.c_i32p), self.copy_beads.ctypes.data_as(_abi.c_i32p),
            C.byref(self._p)), "igmk_pop_create")

    def close(self):
        if self._p:
            self._lib.igmk_pop_destroy(self._p)
            self._p = C.c_void_p()

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
        return False

    # ---- coordinates ------------------------------------------------------------------
    def _check_crd(self, crd):
        if crd.dtype != np.float32 or crd.shape != (self.n_bead, self.n_struct, 3) or not crd.flags.c_contiguous:
            raise ValueError("coordinates must be C-contiguous float32 [%d][%d][3]" % (self.n_bead, self.n_struct))

    def upload(self, crd):
        self._check_crd(crd)
        _abi.check(self._lib.igmk_pop_upload(self._p, crd.ctypes.data_as(_abi.c_f32p)), "igmk_pop_upload")

    def download(self):
        crd = np.empty((self.n_bead, self.n_struct, 3), dtype=np.float32)
        _abi.check(self._lib.igmk_pop_download(self._p, crd.ctypes.data_as(_abi.c_f32p)), "igmk_pop_download")
        return crd

    def attach(self, tensor):
        """Use a caller-owned CUDA tensor (float32 [n_bead][n_struct][3], contiguous) as the slab."""
        if tuple(tensor.shape) != (self.n_bead, self.n_struct, 3) or not tensor.is_contiguous() \
                or str(tensor.dtype) != "torch.float32" or not tensor.is_cuda:
            raise ValueError("attach() needs a contiguous CUDA float32 tensor [%d][%d][3]" % (self.n_bead, self.n_struct))
        _abi.check(self._lib.igmk_pop_attach_coords(self._p, C.c_void_p(tensor.data_ptr())), "igmk_pop_attach_coords")
        self._keep = tensor

    def last_kernel_ms(self):
        return float(self._lib.igmk_pop_last_kernel_ms(self._p))

    # ---- activation distances -----------------------------------------------------------
    def set_pairs(self, pair_i, pair_j, pwish, plast):
        pi = np.ascontiguousarray(pair_i, dtype=np.int32)
        pj = np.ascontiguousarray(pair_j, dtype=np.int32)
        pw = np.ascontiguousarray(pwish, dtype=np.float64)
        pl = np.ascontiguousarray(plast, dtype=np.float64)
        if not (pi.size == pj.size == pw.size == pl.size):
            raise ValueError("pair arrays differ in length")
        self._n_pairs = int(pi.size)
        _abi.check(self._lib.igmk_pop_set_pairs(
            self._p, pi.size, pi.ctypes.data_as(_abi.c_i32p), pj.ctypes.data_as(_abi.c_i32p),
            pw.ctypes.data_as(_abi.c_f64p), pl.ctypes.data_as(_abi.c_f64p)), "igmk_pop_set_pairs")

    def actdist_launch(self, contact_range=2.0):
        _abi.check(self._lib.igmk_pop_actdist_launch(self._p, float(contact_range)), "igmk_pop_actdist_launch")

    def actdist_collect(self):
        cap = max(1, self._n_pairs * self.max_copies * self.max_copies)
        row = np.empty(cap, dtype=np.int32)
        col = np.empty(cap, dtype=np.int32)
        dist = np.empty(cap, dtype=np.float32)
        prob = np.empty(cap, dtype=np.float32)
        n = C.c_int64(0)
        _abi.check(self._lib.igmk_pop_actdist_collect(
            self._p, cap, row.ctypes.data_as(_abi.c_i32p), col.ctypes.data_as(_abi.c_i32p),
            dist.ctypes.data_as(_abi.c_f32p), prob.ctypes.data_as(_abi.c_f32p), C.byref(n)), "igmk_pop_actdist_collect")
        k = n.value
        return row[:k], col[:k], dist[:k], prob[:k]

    def actdist(self, pair_i, pair_j, pwish, plast, contact_range=2.0):
        """-> (row i4, col i4, dist f4, prob f4): the rows this batch contributes to
        actdist.hdf5 (ActivationDistanceStep.py:277-281), bit for bit."""
        self.set_pairs(pair_i, pair_j, pwish, plast)
        self.actdist_launch(contact_range)
        return self.actdist_collect()

    def contact_freq(self, pair_i, pair_j, contact_range=2.0):
        """-> float32 [n_pairs]: fraction of (copy pair, structure) combinations in contact, the
        ``pnow`` of get_actdist (ActivationDistanceStep.py:419-422)."""
        pi = np.ascontiguousarray(pair_i, dtype=np.int32)
        pj = np.ascontiguousarray(pair_j, dtype=np.int32)
        if pi.size != pj.size:
            raise ValueError("pair arrays differ in length")
        out = np.empty(pi.size, dtype=np.float32)
        _abi.check(self._lib.igmk_pop_contact_freq(self._p, pi.size, pi.ctypes.data_as(_abi.c_i32p),
                                                   pj.ctypes.data_as(_abi.c_i32p), float(contact_range),
                                                   out.ctypes.data_as(_abi.c_f32p)), "igmk_pop_contact_freq")
        return out

    # ---- activation distances with the population sharded by structure ------------------------
    def shard_count(self, contact_range, d_counts):
        """local contact count of every staged pair -> device buffer u32 [n_pairs] (its address: ``d_counts``)"""
        _abi.check(self._lib.igmk_pop_shard_count(self._p, float(contact_range), C.c_void_p(int(d_counts))), "igmk_pop_shard_count")

    def shard_rank(self, n_struct_total, d_counts_total, d_want, d_prefix):
        _abi.check(self._lib.igmk_pop_shard_rank(self._p, int(n_struct_total), C.c_void_p(int(d_counts_total)),
                                                 C.c_void_p(int(d_want)), C.c_void_p(int(d_prefix))), "igmk_pop_shard_rank")

    def shard_hist(self, radix_pass, d_prefix, d_hist):
        _abi.check(self._lib.igmk_pop_shard_hist(self._p, int(radix_pass), C.c_void_p(int(d_prefix)), C.c_void_p(int(d_hist))),
                   "igmk_pop_shard_hist")

    def shard_pick(self, radix_pass, d_hist_total, d_want, d_prefix):
        _abi.check(self._lib.igmk_pop_shard_pick(self._p, int(radix_pass), C.c_void_p(int(d_hist_total)), C.c_void_p(int(d_want)),
                                                 C.c_void_p(int(d_prefix))), "igmk_pop_shard_pick")

    # ---- contact assignment ---------------------------------------------------------------
    def set_rows(self, row, col, dist):
        r = np.ascontiguousarray(row, dtype=np.int32)
        c = np.ascontiguousarray(col, dtype=np.int32)
        d = np.ascontiguousarray(dist, dtype=np.float32)
        if not (r.size == c.size == d.size):
            raise ValueError("row arrays differ in length")
        self._n_rows = int(r.size)
        _abi.check(self._lib.igmk_pop_set_rows(self._p, r.size, r.ctypes.data_as(_abi.c_i32p),
                                               c.ctypes.data_as(_abi.c_i32p), d.ctypes.data_as(_abi.c_f32p)),
                   "igmk_pop_set_rows")

    def assign_launch(self):
        _abi.check(self._lib.igmk_pop_assign_launch(self._p), "igmk_pop_assign_launch")

    def set_row_bounds(self, row_r0, row_k):
        """bound distance of every staged row (contact_range * (r_i + r_j), inter_hic.py:55) and the spring constant,
        for the device-side hand-off to the M-step (MStepBatch.submit_population_device)"""
        r0 = np.ascontiguousarray(row_r0, dtype=np.float64)
        _abi.check(self._lib.igmk_pop_set_row_bounds(self._p, r0.size, r0.ctypes.data_as(_abi.c_f64p), float(row_k)),
                   "igmk_pop_set_row_bounds")

    def assign_collect(self):
        nw = (self.n_struct + 31) // 32
        bits = np.zeros((self._n_rows, nw), dtype=np.uint32)
        _abi.check(self._lib.igmk_pop_assign_collect(self._p, bits.ctypes.data_as(_abi.c_u32p)), "igmk_pop_assign_collect")
        return bits

    def assign(self, row, col, dist):
        """-> uint32 [n_rows][ceil(n_struct/32)]; bit s%32 of word s//32 set <=> structure s
        gets a HarmonicUpperBound for that row (inter_hic.py:53-62, intra_hic.py:54-63)."""
        self.set_rows(row, col, dist)
        self.assign_launch()
        return self.assign_collect()


def _damid_actdist(self, loc, pwish, plast, contact_range=0.05, shape="sphere", nucleus_param=5000.0):
    """get_damid_actdist (igm/steps/DamidActivationDistanceStep.py:285-356) for a batch of loci
    -> (bead i4, dist f4, prob f4) exactly as the rows land in damid_actdist.hdf5."""
    lc = np.ascontiguousarray(loc, dtype=np.int32)
    pw = np.ascontiguousarray(pwish, dtype=np.float64)
    pl = np.ascontiguousarray(plast, dtype=np.float64)
    if not (lc.size == pw.size == pl.size):
        raise ValueError("locus arrays differ in length")
    if shape not in ("sphere", "ellipsoid"):
        raise NotImplementedError("shape %s has not been implemented yet." % shape)      # DamidActivationDistanceStep.py:183
    param = np.ascontiguousarray(np.atleast_1d(nucleus_param), dtype=np.float64)
    if param.size != (1 if shape == "sphere" else 3):
        raise ValueError("nucleus_param does not match the shape")
    cap = max(1, lc.size * self.max_copies)
    bead = np.empty(cap, dtype=np.int32)
    dist = np.empty(cap, dtype=np.float32)
    prob = np.empty(cap, dtype=np.float32)
    n = C.c_int64(0)
    _abi.check(self._lib.igmk_pop_damid_actdist(
        self._p, lc.size, lc.ctypes.data_as(_abi.c_i32p), pw.ctypes.data_as(_abi.c_f64p), pl.ctypes.data_as(_abi.c_f64p),
        float(contact_range), 0 if shape == "sphere" else 1, param.ctypes.data_as(_abi.c_f64p), cap,
        bead.ctypes.data_as(_abi.c_i32p), dist.ctypes.data_as(_abi.c_f32p), prob.ctypes.data_as(_abi.c_f32p), C.byref(n)),
        "igmk_pop_damid_actdist")
    return bead[:n.value], dist[:n.value], prob[:n.value]


def _damid_assign(self, bead, dist, contact_range=0.05, semiaxes=(5000.0, 5000.0, 5000.0)):
    """Damid._apply_envelope's predicate (igm/restraints/damid.py:116-130) for all structures:
    bit s of row r set <=> bead[r] of structure s joins the lamina envelope."""
    b = np.ascontiguousarray(bead, dtype=np.int32)
    d = np.ascontiguousarray(dist, dtype=np.float32)
    if b.size != d.size:
        raise ValueError("row arrays differ in length")
    abc = np.ascontiguousarray(semiaxes, dtype=np.float64)
    bits = np.zeros((b.size, (self.n_struct + 31) // 32), dtype=np.uint32)
    _abi.check(self._lib.igmk_pop_damid_assign(self._p, b.size, b.ctypes.data_as(_abi.c_i32p), d.ctypes.data_as(_abi.c_f32p),
                                               float(contact_range), abc.ctypes.data_as(_abi.c_f64p),
                                               bits.ctypes.data_as(_abi.c_u32p)), "igmk_pop_damid_assign")
    return bits


def _fish_ranks(self, hap_i, hap_j=None):
    """get_pair_dists / get_rad_dists + get_min_max_and_idx (igm/steps/FishAssignmentStep.py:31-88)
    for a batch: pairs (hap_i, hap_j) or, with hap_j None / negative entries, radial probes.
    -> (mindists f4 [n][S], maxdists f4 [n][S], idxmin i4 [n][S], idxmax i4 [n][S])"""
    hi = np.ascontiguousarray(hap_i, dtype=np.int32)
    hj = np.full(hi.size, -1, dtype=np.int32) if hap_j is None else np.ascontiguousarray(hap_j, dtype=np.int32)
    if hi.size != hj.size:
        raise ValueError("locus arrays differ in length")
    shape = (hi.size, self.n_struct)
    lo, hi2 = np.empty(shape, dtype=np.float32), np.empty(shape, dtype=np.float32)
    ra, rb = np.empty(shape, dtype=np.int32), np.empty(shape, dtype=np.int32)
    _abi.check(self._lib.igmk_pop_fish_ranks(self._p, hi.size, hi.ctypes.data_as(_abi.c_i32p), hj.ctypes.data_as(_abi.c_i32p),
                                             lo.ctypes.data_as(_abi.c_f32p), hi2.ctypes.data_as(_abi.c_f32p),
                                             ra.ctypes.data_as(_abi.c_i32p), rb.ctypes.data_as(_abi.c_i32p)), "igmk_pop_fish_ranks")
    return lo, hi2, ra, rb


def _sprite_rg2(self, clusters):
    """get_rgs2 (igm/cython_compiled/sprite.pyx:36-101) for many clusters at once.
    clusters: list of clusters, each a list of regions, each a list of alternative bead ids.
    -> list of (rg2s f4 [S], best_structure int, copy_idxs i4 [S][n_regions]) like the reference returns."""
    cl_off, rg_off, beads = [0], [0], []
    for cl in clusters:
        for region in cl:
            beads.extend(int(b) for b in region)
            rg_off.append(len(beads))
        cl_off.append(len(rg_off) - 1)
    cl = np.array(cl_off, dtype=np.int32)
    rg = np.array(rg_off, dtype=np.int32)
    bd = np.array(beads if beads else [0], dtype=np.int32)
    S = self.n_struct
    rg2 = np.empty((len(clusters), S), dtype=np.float32)
    ci = np.empty((max(len(rg_off) - 1, 1), S), dtype=np.int8)
    ms = np.empty(max(len(clusters), 1), dtype=np.int32)
    _abi.check(self._lib.igmk_pop_sprite_rg2(self._p, len(clusters), cl.ctypes.data_as(_abi.c_i32p), rg.ctypes.data_as(_abi.c_i32p),
                                             bd.ctypes.data_as(_abi.c_i32p), rg2.ctypes.data_as(_abi.c_f32p),
                                             ci.ctypes.data_as(C.POINTER(C.c_int8)), ms.ctypes.data_as(_abi.c_i32p)),
               "igmk_pop_sprite_rg2")
    return [(rg2[c], int(ms[c]), np.ascontiguousarray(ci[cl_off[c]:cl_off[c + 1]].T.astype(np.int32))) for c in range(len(clusters))]


Population.sprite_rg2 = _sprite_rg2
Population.fish_ranks = _fish_ranks
Population.damid_actdist = _damid_actdist
Population.damid_assign = _damid_assign


def unpack_bits(bits, n_struct):
    """uint32 [n_rows][nw] -> bool [n_rows][n_struct]"""
    b = (bits[:, :, None] >> np.arange(32, dtype=np.uint32)) & np.uint32(1)
    return b.reshape(bits.shape[0], -1)[:, :n_struct].astype(bool)
