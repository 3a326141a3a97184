"""The ``optimization/kernel = "b200"`` plugin: same contract as the reference's
``igm.model.kernel.lammps.optimize(model, cfg)`` (igm/model/kernel/lammps.py:361-489).

* reads ``cfg['optimization']['optimizer_options']`` (+ ``kernel_opts/b200``),
  ``cfg['runtime']['step_no']`` for the seed (lammps.py:435, :159-165);
* mutates ``model.particles[i].pos`` in place for EVERY particle (lammps.py:464-465);
* returns the ``info`` dict of ``get_info_from_log`` (lammps_io.py:6-37):
  ``final-energy``, ``pair-energy``, ``bond-energy``, ``thermo`` {Temp, E_pair, E_bond,
  f_envelope<j>}, ``md-time``;
* raises ``RuntimeError`` naming the model id when the kernel fails (lammps.py:450-454).

``optimize_population`` is the same thing for many models in ONE launch sequence (a CTA per
structure); ``optimize`` is the batch-of-one case.  No temp files, no subprocess.  There is no
CPU fallback: without the CUDA library or a GPU these functions raise.
"""
import ctypes as C

import numpy as np

from . import _abi
from .pack import pack_model, script_seed

KERNEL_NAME = "b200"          # value of cfg['optimization']['kernel'] that selects this module

_DEFAULT_KERNEL_OPTS = {
    "device": 0,              # CUDA device ordinal
    "bond_prefactor": 1.0,    # E = bond_prefactor * k * (r - r0)^2 (LAMMPS `harmonic` convention)
}


def _cfg_get(cfg, path, default=None):
    """cfg may be an igm.core.Config (slash paths, config.py:98-113) or a plain dict."""
    node = cfg
    for key in path.split("/"):
        try:
            node = node[key]
        except (KeyError, TypeError, IndexError):
            return default
    return node


def kernel_opts(cfg):
    opts = dict(_DEFAULT_KERNEL_OPTS)
    opts.update(_cfg_get(cfg, "optimization/kernel_opts/" + KERNEL_NAME, {}) or {})
    return opts


STAGING_ONLY = -1       # MStepBatch(STAGING_ONLY): host staging without a GPU (submit*, slot_sizes, layout_stats); no compute


class MStepBatch(object):
    """Thin owner of one ``igmk_handle_t`` (one device)."""

    def __init__(self, device=0):
        self._lib = _abi.load()
        self._h = C.c_void_p()
        _abi.check(self._lib.igmk_create(int(device), C.byref(self._h)), "igmk_create")
        self._packed = []

    def close(self):
        if self._h:
            self._lib.igmk_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
        return False

    def __len__(self):
        return len(self._packed)

    def clear(self):
        _abi.check(self._lib.igmk_clear(self._h), "igmk_clear")
        self._packed = []

    def submit(self, packed):
        slot = C.c_int32(-1)
        _abi.check(self._lib.igmk_submit(self._h, C.byref(packed.c), C.byref(slot)),
                   "igmk_submit (modelid: %d)" % packed.struct_id)
        self._packed.append(packed)
        return slot.value

    def submit_population(self, template, crd, n_bead, row, col, row_r0, row_k, bits, s_begin=0, n_struct=None, seeds=None,
                          bits_word0=0, extras=None):
        """Submit structures ``s_begin .. s_begin+n_struct-1`` of the .hss-layout slab ``crd``
        ([n_bead][S][3] float32): every structure = ``template`` (a PackedSystem: shared particles and
        restraints) with the slab's coordinates and a HarmonicUpperBound per actdist row whose bit is set
        in ``bits`` (the matrix Population.assign returned for exactly these structures; ``bits_word0`` =
        word of every row that holds the bit of structure ``s_begin``, for a matrix that covers more
        structures than this call submits).  ``extras``: a PopulationExtras with the restraints that differ by
        structure beyond the Hi-C rows (FISH bonds, DamID envelope members); it covers the same structures as
        ``crd`` (global indices).  -> first slot."""
        crd = np.ascontiguousarray(crd, dtype=np.float32)
        S = crd.shape[1]
        if n_struct is None:
            n_struct = S - s_begin
        r = np.ascontiguousarray(row, dtype=np.int32)
        c = np.ascontiguousarray(col, dtype=np.int32)
        r0 = np.ascontiguousarray(row_r0, dtype=np.float64)
        b = np.ascontiguousarray(bits, dtype=np.uint32).reshape(r.size, -1) if r.size else np.zeros((0, 1), dtype=np.uint32)
        if r.size and (bits_word0 < 0 or (b.shape[1] - bits_word0) * 32 < n_struct):
            raise ValueError("bit matrix does not cover structures %d..%d from word %d" % (s_begin, s_begin + n_struct, bits_word0))
        sd = None if seeds is None else np.ascontiguousarray(seeds, dtype=np.int64)
        first = C.c_int32(-1)
        ex = None if extras is None else extras.c_view(s_begin, n_struct)
        _abi.check(self._lib.igmk_submit_population_ex(
            self._h, C.byref(template.c), int(n_bead), int(S), int(s_begin), int(n_struct), crd.ctypes.data_as(_abi.c_f32p),
            r.size, r.ctypes.data_as(_abi.c_i32p), c.ctypes.data_as(_abi.c_i32p), r0.ctypes.data_as(_abi.c_f64p), float(row_k),
            C.cast(b.ctypes.data + 4 * int(bits_word0), _abi.c_u32p), int(b.shape[1]),
            None if sd is None else sd.ctypes.data_as(_abi.c_i64p),
            None if ex is None else C.byref(ex[0]),
            C.byref(first)), "igmk_submit_population")
        self._packed.extend([template] * n_struct)        # sizes for fetch(); bond arrays differ per structure
        return first.value

    def submit_population_device(self, template, pop, s_begin=0, n_struct=None, seeds=None):
        """The same for structures of a device-resident population whose assignment is still on the GPU
        (``pop``: an astep.Population after set_rows + assign_launch + set_row_bounds): nothing but the template
        crosses the bus, the batch is assembled by kernels (igmk_submit_population_dev).  -> True, or False when
        the input is outside what the device path covers and has to go through submit_population."""
        if n_struct is None:
            n_struct = pop.n_struct - s_begin
        sd = None if seeds is None else np.ascontiguousarray(seeds, dtype=np.int64)
        first = C.c_int32(-1)
        rc = self._lib.igmk_submit_population_dev(self._h, C.byref(template.c), pop._p, int(s_begin), int(n_struct),
                                                  None if sd is None else sd.ctypes.data_as(_abi.c_i64p), C.byref(first))
        if rc == 2:
            return False
        _abi.check(rc, "igmk_submit_population_dev")
        self._packed.extend([template] * n_struct)
        return True

    def debug_array(self, which, dtype=np.uint8):
        """array `which` of the staged batch as it sits on the device (igmk_debug_fetch_array)"""
        n = C.c_int64(0)
        _abi.check(self._lib.igmk_debug_fetch_array(self._h, int(which), None, 0, C.byref(n)), "igmk_debug_fetch_array")
        buf = np.zeros(max(n.value, 1), dtype=np.uint8)
        _abi.check(self._lib.igmk_debug_fetch_array(self._h, int(which), buf.ctypes.data_as(C.c_void_p), buf.size, C.byref(n)),
                   "igmk_debug_fetch_array")
        return buf[:n.value].view(dtype)

    def upload(self):
        _abi.check(self._lib.igmk_upload(self._h), "igmk_upload")

    def reset_coordinates(self):
        _abi.check(self._lib.igmk_reset_coordinates(self._h), "igmk_reset_coordinates")

    def run(self, protocol):
        _abi.check(self._lib.igmk_run(self._h, C.byref(protocol)), "igmk_run")

    def single_point(self, ev_scale=1.0, env_scale=1.0, bond_prefactor=1.0, f32=False):
        fn = self._lib.igmk_single_point_f32 if f32 else self._lib.igmk_single_point
        _abi.check(fn(self._h, float(ev_scale), float(env_scale), float(bond_prefactor)), "igmk_single_point")

    def md_forces(self, protocol, stage=0):
        """forces of the MD kernel's own (float32) force routine at the submitted coordinates, run setup of `stage`;
        read them with fetch_forces"""
        _abi.check(self._lib.igmk_md_forces(self._h, C.byref(protocol), int(stage)), "igmk_md_forces")

    def layout_stats(self):
        """-> (slots per thread K, bond items, items after sliced-ELL padding, host seconds of the layout build)"""
        k, b, e, t = C.c_int32(0), C.c_int64(0), C.c_int64(0), C.c_double(0.0)
        _abi.check(self._lib.igmk_layout_stats(self._h, C.byref(k), C.byref(b), C.byref(e), C.byref(t)), "igmk_layout_stats")
        return k.value, b.value, e.value, t.value

    def slot_sizes(self, slot):
        """-> (n_atoms, n_bonds, n_envelope_members) of a submitted structure"""
        a, b, e = C.c_int32(0), C.c_int64(0), C.c_int64(0)
        _abi.check(self._lib.igmk_slot_sizes(self._h, slot, C.byref(a), C.byref(b), C.byref(e)), "igmk_slot_sizes")
        return a.value, b.value, e.value

    def fetch(self, slot):
        n = self.slot_sizes(slot)[0]
        xyz = np.empty((n, 3), dtype=np.float32)
        res = _abi.Result()
        _abi.check(self._lib.igmk_fetch(self._h, slot, xyz.ctypes.data_as(_abi.c_f32p), C.byref(res)), "igmk_fetch")
        return xyz, res

    def fetch_population(self, crd, s_begin=0):
        """Write every structure of the batch into the slab ``crd`` (float32 [n_bead][S][3], C-contiguous, in
        place) at structures ``s_begin ..``; -> ctypes array of Result, one per structure."""
        if crd.dtype != np.float32 or not crd.flags.c_contiguous or crd.ndim != 3 or crd.shape[2] != 3:
            raise ValueError("crd must be a C-contiguous float32 [n_bead][S][3] array")
        res = (_abi.Result * max(len(self), 1))()
        _abi.check(self._lib.igmk_fetch_population(self._h, crd.shape[0], crd.shape[1], int(s_begin),
                                                   crd.ctypes.data_as(_abi.c_f32p), res), "igmk_fetch_population")
        return res

    def violation_ratios(self, slot):
        """-> (bond_ratio f8 [n_bonds] in submit order, env_u f8 [sum of envelope members]) at the
        current coordinates; see igmk_violation_ratios in include/igmk.h."""
        _, nb, ne = self.slot_sizes(slot)
        br = np.zeros(nb, dtype=np.float64)
        eu = np.zeros(ne, dtype=np.float64)
        _abi.check(self._lib.igmk_violation_ratios(self._h, slot, br.ctypes.data_as(_abi.c_f64p), eu.ctypes.data_as(_abi.c_f64p)),
                   "igmk_violation_ratios")
        return br, eu

    def last_timings(self):
        """-> (md_ms, cg_ms, h2d_bytes) of the last run() / upload()"""
        a, b, c = C.c_double(0), C.c_double(0), C.c_int64(0)
        _abi.check(self._lib.igmk_last_timings(self._h, C.byref(a), C.byref(b), C.byref(c)), "igmk_last_timings")
        return a.value, b.value, c.value

    def fetch_forces(self, slot):
        n = self.slot_sizes(slot)[0]
        f = np.empty((n, 3), dtype=np.float64)
        _abi.check(self._lib.igmk_fetch_forces(self._h, slot, f.ctypes.data_as(_abi.c_f64p)), "igmk_fetch_forces")
        return f


class PopulationExtras(object):
    """Restraints of a population that differ from structure to structure beyond the assigned Hi-C rows
    (igmk_pop_extras_t): per-structure bonds in CSR form and envelope groups whose members are given by a bit
    matrix over the structures.

    bonds: (off i8 [S+1], i i4, j i4, k f8, r0 f8, style u1) or None -- structure s owns entries off[s]..off[s+1];
    envelopes: list of (abck (a, b, c, k), bead ids i4 [n_rows], bits u4 [n_rows][ceil(S/32)]) -- bit s of row r set
    means bead ids[r] belongs to the group in structure s (the matrix Population.damid_assign returns);
    atoms: (off i8 [S+1], xyz, radius, kind) -- particles a structure has on top of the template's (SPRITE centroids)."""

    def __init__(self, n_struct, bonds=None, envelopes=(), atoms=None):
        self.n_struct = int(n_struct)
        if atoms is not None:                      # (off i8 [S+1], xyz f4 [n][3], radius f4, kind u1): a structure's own particles
            off, xyz, rad, kind = atoms
            self.atom_off = np.ascontiguousarray(off, dtype=np.int64)
            self.atom_xyz = np.ascontiguousarray(xyz, dtype=np.float32).reshape(-1, 3)
            self.atom_radius = np.ascontiguousarray(rad, dtype=np.float32)
            self.atom_kind = np.ascontiguousarray(kind, dtype=np.uint8)
            if self.atom_off.size != self.n_struct + 1 or not (self.atom_xyz.shape[0] == self.atom_radius.size ==
                                                                 self.atom_kind.size == int(self.atom_off[-1])):
                raise ValueError("per-structure atom arrays are inconsistent")
        else:
            self.atom_off = None
        if bonds is not None:
            off, bi, bj, bk, br0, bs = bonds
            self.bond_off = np.ascontiguousarray(off, dtype=np.int64)
            self.bond_i = np.ascontiguousarray(bi, dtype=np.int32)
            self.bond_j = np.ascontiguousarray(bj, dtype=np.int32)
            self.bond_k = np.ascontiguousarray(bk, dtype=np.float64)
            self.bond_r0 = np.ascontiguousarray(br0, dtype=np.float64)
            self.bond_style = np.ascontiguousarray(bs, dtype=np.uint8)
            if self.bond_off.size != self.n_struct + 1 or not (self.bond_i.size == self.bond_j.size == self.bond_k.size ==
                                                                 self.bond_r0.size == self.bond_style.size == int(self.bond_off[-1])):
                raise ValueError("per-structure bond arrays are inconsistent")
        else:
            self.bond_off = None
        nw = (self.n_struct + 31) // 32
        self.env_abck = np.ascontiguousarray([e[0] for e in envelopes], dtype=np.float64).reshape(-1, 4)
        beads = [np.ascontiguousarray(e[1], dtype=np.int32).ravel() for e in envelopes]
        bits = [np.ascontiguousarray(e[2], dtype=np.uint32).reshape(b.size, nw) for e, b in zip(envelopes, beads)]
        self.env_row_off = np.concatenate([[0], np.cumsum([b.size for b in beads])]).astype(np.int64)
        self.env_bead = np.concatenate(beads) if beads else np.zeros(0, dtype=np.int32)
        self.env_bits = np.concatenate(bits) if bits else np.zeros((0, nw), dtype=np.uint32)
        self._nw = nw

    def c_view(self, s_begin, n_struct):
        """-> (PopExtras, keep-alive list) for structures s_begin .. s_begin+n_struct-1 (s_begin a multiple of 32
        when there are envelope bits)"""
        if s_begin < 0 or s_begin + n_struct > self.n_struct:
            raise ValueError("structure range outside the extras")
        ex = _abi.PopExtras()
        keep = []
        if self.bond_off is not None:
            off = np.ascontiguousarray(self.bond_off[s_begin:s_begin + n_struct + 1])       # absolute offsets into the full arrays
            keep.append(off)
            ex.bond_off = off.ctypes.data_as(_abi.c_i64p)
            ex.bond_i = self.bond_i.ctypes.data_as(_abi.c_i32p)
            ex.bond_j = self.bond_j.ctypes.data_as(_abi.c_i32p)
            ex.bond_k = self.bond_k.ctypes.data_as(_abi.c_f64p)
            ex.bond_r0 = self.bond_r0.ctypes.data_as(_abi.c_f64p)
            ex.bond_style = self.bond_style.ctypes.data_as(_abi.c_u8p)
        if self.atom_off is not None:
            aoff = np.ascontiguousarray(self.atom_off[s_begin:s_begin + n_struct + 1])
            keep.append(aoff)
            ex.atom_off = aoff.ctypes.data_as(_abi.c_i64p)
            ex.atom_xyz = self.atom_xyz.ctypes.data_as(_abi.c_f32p)
            ex.atom_radius = self.atom_radius.ctypes.data_as(_abi.c_f32p)
            ex.atom_kind = self.atom_kind.ctypes.data_as(_abi.c_u8p)
        ex.n_env = int(self.env_abck.shape[0])
        if ex.n_env:
            if s_begin % 32:
                raise ValueError("with per-structure envelopes a chunk must start at a multiple of 32 structures")
            ex.env_abck = self.env_abck.ctypes.data_as(_abi.c_f64p)
            ex.env_row_off = self.env_row_off.ctypes.data_as(_abi.c_i64p)
            ex.env_bead = self.env_bead.ctypes.data_as(_abi.c_i32p)
            ex.env_bits = C.cast(self.env_bits.ctypes.data + 4 * (s_begin // 32), _abi.c_u32p)
            ex.env_bits_words = self._nw
        return ex, keep


def merge_bond_lists(n_struct, *lists):
    """several per-structure bond lists (off, i, j, k, r0, style) -> one, structure by structure in the order given
    (ModelingStep.task adds Sprite before Fish, igm/steps/ModelingStep.py:285-330)"""
    lists = [b for b in lists if b is not None]
    if not lists:
        return None
    if len(lists) == 1:
        return lists[0]
    counts = np.sum([np.diff(b[0]) for b in lists], axis=0)
    off = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    out = [np.empty(int(off[-1]), dtype=np.asarray(lists[0][c]).dtype) for c in range(1, 6)]
    cur = off[:-1].copy()
    for b in lists:
        n = np.diff(b[0])
        dst = (np.repeat(cur, n) + (np.arange(int(n.sum())) - np.repeat(b[0][:-1], n))).astype(np.int64)
        for c in range(5):
            out[c][dst] = np.asarray(b[c + 1])
        cur = cur + n
    return (off,) + tuple(out)


def info_from_result(res):
    """igmk_result_t -> the dict get_info_from_log builds (lammps_io.py:6-37)."""
    thermo = {"Temp": res.temperature, "E_pair": res.pair_energy, "E_bond": res.bond_energy}
    for j in range(res.n_env):
        thermo["f_envelope%d" % j] = res.env_energy[j]        # fix ids, lammps.py:296
    return {
        "final-energy": res.final_energy,
        "pair-energy": res.pair_energy,
        "bond-energy": res.bond_energy,
        "thermo": thermo,
        "md-time": res.md_seconds,
    }


def envelope_ratios(env_u, k, scale):
    """Finish EllipticEnvelope.getViolationRatios (igm/model/forces.py:236-250) from the kernel's
    scale-free part: t = u/scale for k > 0 (k2 > 1), t = u for k < 0 (k2 < 1); score = max(0, t*k);
    ratio = score / (k*scale)."""
    t = env_u / scale if k > 0 else env_u
    return np.maximum(0, t * k) / (k * scale)


def violation_histogram(v, nbins=100, vmax=1):
    """get_violation_histogram (igm/steps/ModelingStep.py:649-659), verbatim semantics"""
    v = np.array(v)
    over = np.count_nonzero(v > vmax)
    inner = v[v <= vmax]
    H, edges = np.histogram(inner, bins=nbins, range=(0, vmax))
    return np.concatenate([H, [over]]), np.concatenate([edges, [np.inf]])


PIPELINE_CHUNK = 128     # structures per launch sequence of run_population: a multiple of 32 (bit-matrix words), just under one wave of 148 SMs


def run_population(template, crd, n_bead, row, col, row_r0, row_k, bits, protocol, device=0, seeds=None,
                   chunk=PIPELINE_CHUNK, violations_tol=None, extras=None, pop=None):
    """M-step of a whole assigned population, host buffers in, host buffers out: the population-wide
    form of ``ModelingStep.task`` (igm/steps/ModelingStep.py:184-264 build, :266 optimize, :341-382 results).

    The structures go to the GPU in chunks of ``chunk``: this thread assembles chunk c+1
    (``igmk_submit_population``, multi-threaded C++) while worker threads upload, run and fetch the
    chunks before it, each batch on its own stream -- the SMs always have CTAs of some chunk to run, so
    the host-side assembly, the copies in and the copies out disappear behind the MD kernel.

    ``pop``: the astep.Population these structures and rows were assigned in (set_rows + assign_launch done on it,
    same structure indices as ``crd``).  Then the batches are assembled on the device from the population's slab, rows
    and bit matrix (igmk_submit_population_dev) -- ``crd`` and ``bits`` are not read, ``bits`` may be None -- unless
    the input is outside what that path covers (per-structure extras, very many distinct bounds), in which case the
    bit matrix is fetched and the host path runs.

    -> (coordinates float32 [n_bead][S][3], [Result] * S, violations or None, timings dict);
    violations = per-structure (n_imposed, n_violated) of the harmonic restraints and (n_imposed, n_violated) of the
    envelope restraints (one per member, scale = 0.1 * mean semi-axis, envelope.py:51) at ``violations_tol``."""
    import threading
    crd = np.ascontiguousarray(crd, dtype=np.float32)
    S = crd.shape[1]
    chunk = max(32, (int(chunk) + 31) // 32 * 32)
    new = np.empty((n_bead, S, 3), dtype=np.float32)
    results, viol = [None] * S, ([None] * S if violations_tol is not None else None)
    errors, workers = [], []
    tim = {"md_kernel_ms": 0.0, "cg_kernel_ms": 0.0, "h2d_bytes": 0, "chunks": 0}
    lock = threading.Lock()

    def finish(batch, lo, hi):
        try:
            batch.run(protocol)
            res = batch.fetch_population(new, lo)
            for k in range(hi - lo):
                results[lo + k] = res[k]
            md, cg, h2d = batch.last_timings()
            with lock:
                tim["md_kernel_ms"] += md; tim["cg_kernel_ms"] += cg; tim["h2d_bytes"] += h2d; tim["chunks"] += 1
        except Exception as e:           # re-raised on the calling thread
            errors.append(e)

    # The batches stay open until every chunk is done: freeing device memory waits for the whole device
    # (and holds up the other threads' launches while it does), which would stall the pipeline.
    batches = []
    tim["assembled_on_device"] = 0
    try:
        if pop is not None and extras is None:
            # Assembly on the device costs milliseconds and its kernels would queue behind the MD CTAs of the chunks before
            # them (an MD CTA fills its SM), so there is nothing to pipeline: one batch, one launch sequence.
            pop.set_row_bounds(row_r0, row_k)
            batches.append((MStepBatch(device), 0, S))
            if batches[0][0].submit_population_device(template, pop, 0, S, seeds=seeds):
                tim["assembled_on_device"] = 1
                finish(*batches[0])
                if errors:
                    raise errors[0]
            else:
                batches.pop()[0].close()
        if not tim["assembled_on_device"] and bits is None:
            if pop is None:
                raise ValueError("run_population needs the assignment bit matrix or the population it was computed in")
            bits = pop.assign_collect()
        for lo in ([] if tim["assembled_on_device"] else range(0, S, chunk)):
            hi = min(S, lo + chunk)
            batches.append((MStepBatch(device), lo, hi))
            batches[-1][0].submit_population(template, crd, n_bead, row, col, row_r0, row_k, bits, s_begin=lo, n_struct=hi - lo,
                                             seeds=None if seeds is None else seeds[lo:hi], bits_word0=lo // 32, extras=extras)
            w = threading.Thread(target=finish, args=batches[-1])
            w.start()
            workers.append(w)
            if errors:
                break
        for w in workers:
            w.join()
        workers = []
        if errors:
            raise errors[0]
        if viol is not None:
            for batch, lo, hi in batches:
                for k in range(hi - lo):
                    br, eu = batch.violation_ratios(k)
                    n_env = n_env_vio = 0
                    for e in range(template.env_abck.shape[0]):      # Envelope restraints: EllipticEnvelope per member
                        a, b, c, ke = (float(v) for v in template.env_abck[e])
                        u = eu[template.env_off[e]:template.env_off[e + 1]]
                        n_env += u.size
                        if ke != 0.0:
                            n_env_vio += int(np.count_nonzero(envelope_ratios(u, ke, 0.1 * (a + b + c) / 3.0) > violations_tol))   # envelope.py:51
                    viol[lo + k] = (br.size, int(np.count_nonzero(br > violations_tol)), n_env, n_env_vio)
    finally:
        for w in workers:
            w.join()
        for batch, _, _ in batches:
            batch.close()
    return new, results, viol, tim


def _optimize_on_device(models, seeds, protocol, device, chunk=PIPELINE_CHUNK):
    """Pack and submit ``chunk`` models at a time on this thread; a worker thread per chunk runs it and
    writes the positions back, so packing (Python) overlaps the MD of the chunks before it."""
    import threading
    infos, errors, workers = [None] * len(models), [], []

    def finish(batch, lo, slots):
        try:
            batch.run(protocol)
            for k, slot in enumerate(slots):
                m = models[lo + k]
                xyz, res = batch.fetch(slot)
                if res.status != 0 or not np.isfinite(res.final_energy):
                    raise RuntimeError("b200 kernel produced non-finite coordinates or energy (modelid: %d)" % m.id)
                for i, p in enumerate(m.particles):
                    p.pos = xyz[i]                                                # lammps.py:464-465
                infos[lo + k] = info_from_result(res)
        except Exception as e:           # re-raised on the calling thread
            errors.append(e)

    batches = []                         # closed together at the end (see run_population)
    try:
        for lo in range(0, len(models), chunk):
            hi = min(len(models), lo + chunk)
            batch = MStepBatch(device)
            batches.append(batch)
            slots = [batch.submit(pack_model(m, seed=sd)) for m, sd in zip(models[lo:hi], seeds[lo:hi])]
            w = threading.Thread(target=finish, args=(batch, lo, slots))
            w.start()
            workers.append(w)
            if errors:
                break
    finally:
        for w in workers:
            w.join()
        for batch in batches:
            batch.close()
    if errors:
        raise errors[0]
    return infos


def _devices(kopts, device):
    if device is not None:
        return [int(device)]
    devs = kopts.get("devices", None)
    if devs is None:
        return [int(kopts["device"])]
    if devs == "all":
        return list(range(_abi.load().igmk_device_count()))
    return [int(d) for d in devs]


def optimize_population(models, cfg, device=None):
    """Optimise every model of ``models`` in one launch sequence per GPU; returns the list of info dicts.
    ``kernel_opts/b200/devices`` (a list of ordinals or "all") spreads the structures over several GPUs of
    the box from this one process: contiguous blocks, one host thread and one batch per device, no
    communication (structures are independent, igm/steps/ModelingStep.py:97)."""
    models = list(models)
    if not models:
        return []
    kopts = kernel_opts(cfg)
    run_opts = dict(_cfg_get(cfg, "optimization/optimizer_options"))
    run_opts.update(kopts)                                                        # lammps.py:433 (carries `seed`)
    step_no = (_cfg_get(cfg, "runtime/step_no", 1) or 0) + 2                     # lammps.py:435
    protocol = _abi.make_protocol(run_opts, bond_prefactor=kopts["bond_prefactor"])
    seeds = [script_seed(run_opts, m.id, step_no) for m in models]
    devs = _devices(kopts, device)[:len(models)]
    if len(devs) <= 1:
        return _optimize_on_device(models, seeds, protocol, devs[0] if devs else 0)
    from concurrent.futures import ThreadPoolExecutor
    from .parallel import structure_shard
    blocks = [structure_shard(len(models), r, len(devs)) for r in range(len(devs))]
    with ThreadPoolExecutor(max_workers=len(devs)) as ex:
        parts = list(ex.map(lambda rb: _optimize_on_device(models[rb[1][0]:rb[1][1]], seeds[rb[1][0]:rb[1][1]], protocol, devs[rb[0]]),
                            enumerate(blocks)))
    return [info for part in parts for info in part]


def optimize(model, cfg):
    """Drop-in for ``igm.model.kernel.lammps.optimize`` (lammps.py:361).  Under
    ``igm_b200.controller.B200Controller`` concurrent calls are gathered into one launch."""
    from .controller import current_batcher
    batcher = current_batcher()
    if batcher is not None:
        return batcher.optimize(model, cfg)
    return optimize_population([model], cfg)[0]
