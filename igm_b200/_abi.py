"""ctypes mirror of include/igmk.h and the loader of the CUDA library.

There is deliberately no fallback: if ``libigmk.so`` (built in-tree by
``__graft_entry__.build()``) is missing or does not export a declared symbol,
importing the product path raises.
"""
import ctypes as C
import os

import numpy as np

IGMK_MAX_STAGES = 32
IGMK_MAX_ENVELOPES = 8

KIND_BEAD, KIND_STATIC, KIND_CENTROID = 0, 1, 2
BOND_UPPER, BOND_LOWER = 0, 1

CG_STOP = ["max iterations", "max force evaluations", "energy tolerance", "force tolerance",
           "search direction not downhill", "zero force", "zero alpha", "zero quadratic"]

c_f32p = C.POINTER(C.c_float)
c_f64p = C.POINTER(C.c_double)
c_i32p = C.POINTER(C.c_int32)
c_u32p = C.POINTER(C.c_uint32)
c_u8p = C.POINTER(C.c_uint8)
c_i64p = C.POINTER(C.c_int64)


class System(C.Structure):
    _fields_ = [
        ("n_atoms", C.c_int32), ("xyz", c_f32p), ("radius", c_f32p), ("kind", c_u8p),
        ("n_bonds", C.c_int32), ("bond_i", c_i32p), ("bond_j", c_i32p),
        ("bond_k", c_f64p), ("bond_r0", c_f64p), ("bond_style", c_u8p),
        ("n_env", C.c_int32), ("env_abck", c_f64p), ("env_off", c_i32p), ("env_members", c_i32p),
        ("evfactor", C.c_double), ("seed", C.c_int64), ("struct_id", C.c_int64),
    ]


class Protocol(C.Structure):
    _fields_ = [
        ("n_stage", C.c_int32),
        ("mdsteps", C.c_int32 * IGMK_MAX_STAGES),
        ("tstart", C.c_double * IGMK_MAX_STAGES),
        ("tstop", C.c_double * IGMK_MAX_STAGES),
        ("evfactor", C.c_double * IGMK_MAX_STAGES),
        ("envfactor", C.c_double * IGMK_MAX_STAGES),
        ("relax_mdsteps", C.c_int32),
        ("relax_temperature", C.c_double),
        ("relax_max_velocity", C.c_double),
        ("timestep", C.c_double),
        ("max_velocity", C.c_double),
        ("etol", C.c_double), ("ftol", C.c_double),
        ("max_cg_iter", C.c_int32), ("max_cg_eval", C.c_int32),
        ("bond_prefactor", C.c_double),
        ("skip_minimize", C.c_int32),
    ]


class Result(C.Structure):
    _fields_ = [
        ("final_energy", C.c_double), ("pair_energy", C.c_double), ("bond_energy", C.c_double),
        ("temperature", C.c_double), ("env_energy", C.c_double * IGMK_MAX_ENVELOPES),
        ("n_env", C.c_int32), ("cg_iters", C.c_int32), ("cg_evals", C.c_int32), ("cg_stop", C.c_int32),
        ("n_rebuilds", C.c_int32), ("status", C.c_int32), ("md_seconds", C.c_double),
    ]


class PopExtras(C.Structure):
    """igmk_pop_extras_t: restraints that differ from structure to structure (see include/igmk.h)"""
    _fields_ = [
        ("bond_off", c_i64p), ("bond_i", c_i32p), ("bond_j", c_i32p), ("bond_k", c_f64p), ("bond_r0", c_f64p),
        ("bond_style", c_u8p),
        ("n_env", C.c_int32), ("env_abck", c_f64p), ("env_row_off", c_i64p), ("env_bead", c_i32p), ("env_bits", c_u32p),
        ("env_bits_words", C.c_int32),
        ("atom_off", c_i64p), ("atom_xyz", c_f32p), ("atom_radius", c_f32p), ("atom_kind", c_u8p),
    ]


def _ptr(a, ct):
    return a.ctypes.data_as(ct)


class PackedSystem(object):
    """Flat arrays of one structure + the ctypes view that borrows them."""

    def __init__(self, xyz, radius, kind, bond_i, bond_j, bond_k, bond_r0, bond_style,
                 env_abck, env_off, env_members, evfactor, seed=1, struct_id=0):
        self.xyz = np.ascontiguousarray(xyz, dtype=np.float32).reshape(-1, 3)
        n = self.xyz.shape[0]
        self.radius = np.ascontiguousarray(radius, dtype=np.float32).reshape(n)
        self.kind = np.ascontiguousarray(kind, dtype=np.uint8).reshape(n)
        self.bond_i = np.ascontiguousarray(bond_i, dtype=np.int32).ravel()
        self.bond_j = np.ascontiguousarray(bond_j, dtype=np.int32).ravel()
        self.bond_k = np.ascontiguousarray(bond_k, dtype=np.float64).ravel()
        self.bond_r0 = np.ascontiguousarray(bond_r0, dtype=np.float64).ravel()
        self.bond_style = np.ascontiguousarray(bond_style, dtype=np.uint8).ravel()
        nb = self.bond_i.size
        if not (self.bond_j.size == self.bond_k.size == self.bond_r0.size == self.bond_style.size == nb):
            raise ValueError("bond arrays differ in length")
        self.env_abck = np.ascontiguousarray(env_abck, dtype=np.float64).reshape(-1, 4)
        self.env_off = np.ascontiguousarray(env_off, dtype=np.int32).ravel()
        self.env_members = np.ascontiguousarray(env_members, dtype=np.int32).ravel()
        ne = self.env_abck.shape[0]
        if ne > IGMK_MAX_ENVELOPES:
            raise ValueError("at most %d envelopes per structure" % IGMK_MAX_ENVELOPES)
        if self.env_off.size != ne + 1:
            raise ValueError("env_off must have n_env+1 entries")
        self.evfactor = float(evfactor)
        self.seed = int(seed)
        self.struct_id = int(struct_id)
        s = System()
        s.n_atoms = n
        s.xyz = _ptr(self.xyz, c_f32p)
        s.radius = _ptr(self.radius, c_f32p)
        s.kind = _ptr(self.kind, c_u8p)
        s.n_bonds = nb
        s.bond_i = _ptr(self.bond_i, c_i32p)
        s.bond_j = _ptr(self.bond_j, c_i32p)
        s.bond_k = _ptr(self.bond_k, c_f64p)
        s.bond_r0 = _ptr(self.bond_r0, c_f64p)
        s.bond_style = _ptr(self.bond_style, c_u8p)
        s.n_env = ne
        s.env_abck = _ptr(self.env_abck, c_f64p)
        s.env_off = _ptr(self.env_off, c_i32p)
        s.env_members = _ptr(self.env_members, c_i32p)
        s.evfactor = self.evfactor
        s.seed = self.seed
        s.struct_id = self.struct_id
        self.c = s

    @property
    def n_atoms(self):
        return self.xyz.shape[0]


def make_protocol(opts, bond_prefactor=1.0, skip_minimize=False):
    """``optimization/optimizer_options`` dict -> Protocol, following the
    defaults of create_lammps_script (igm/model/kernel/lammps.py:263-283)."""
    prot = opts.get("custom_annealing_protocol", None)
    if prot is None:
        prot = {"num_steps": 1, "mdsteps": [opts["mdsteps"]], "tstarts": [opts["tstart"]],
                "tstops": [opts["tstop"]], "evfactors": [1], "envelope_scales": [1]}
    ns = int(prot.get("num_steps"))
    if ns > IGMK_MAX_STAGES:
        raise ValueError("at most %d annealing stages" % IGMK_MAX_STAGES)
    mdsteps = prot.get("mdsteps", [opts.get("mdsteps")] * ns)
    tstarts = prot.get("tstarts", [opts.get("tstart")] * ns)
    tstops = prot.get("tstops", tstarts)
    evfs = prot.get("evfactors", [1] * ns)
    envfs = prot.get("envelope_factors", [1] * ns)
    if not (len(mdsteps) == len(tstarts) == len(tstops) == len(evfs) == len(envfs) == ns):
        raise AssertionError("annealing protocol lists must all have num_steps entries")
    p = Protocol()
    p.n_stage = ns
    for k in range(ns):
        p.mdsteps[k] = int(mdsteps[k])
        p.tstart[k] = float(tstarts[k])
        p.tstop[k] = float(tstops[k])
        p.evfactor[k] = float(evfs[k])
        p.envfactor[k] = float(envfs[k])
    relax = prot.get("relax", None)
    if relax is not None:
        p.relax_mdsteps = int(relax["mdsteps"])
        p.relax_temperature = float(relax["temperature"])
        p.relax_max_velocity = float(relax["max_velocity"])
    else:
        p.relax_mdsteps = 0
    p.timestep = float(opts["timestep"])
    p.max_velocity = float(opts["max_velocity"])
    p.etol = float(opts["etol"])
    p.ftol = float(opts["ftol"])
    p.max_cg_iter = int(opts["max_cg_iter"])
    p.max_cg_eval = int(opts["max_cg_eval"])
    p.bond_prefactor = float(bond_prefactor)
    p.skip_minimize = 1 if skip_minimize else 0
    return p


# ---------------------------------------------------------------------------------
_LIB = None
_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "libigmk.so")

_vp = C.c_void_p
_SIGNATURES = {
    "igmk_last_error": (C.c_char_p, []),
    "igmk_version": (C.c_int, []),
    "igmk_device_count": (C.c_int, []),
    "igmk_peak_probe": (C.c_int, [C.c_int, c_f64p, c_f64p, c_f64p]),
    "igmk_create": (C.c_int, [C.c_int, C.POINTER(_vp)]),
    "igmk_destroy": (C.c_int, [_vp]),
    "igmk_clear": (C.c_int, [_vp]),
    "igmk_submit": (C.c_int, [_vp, C.POINTER(System), c_i32p]),
    "igmk_submit_population": (C.c_int, [_vp, C.POINTER(System), C.c_int32, C.c_int32, C.c_int32, C.c_int32, c_f32p, C.c_int64,
                                         c_i32p, c_i32p, c_f64p, C.c_double, c_u32p, C.c_int32, c_i64p, c_i32p]),
    "igmk_submit_population_ex": (C.c_int, [_vp, C.POINTER(System), C.c_int32, C.c_int32, C.c_int32, C.c_int32, c_f32p, C.c_int64,
                                            c_i32p, c_i32p, c_f64p, C.c_double, c_u32p, C.c_int32, c_i64p, C.POINTER(PopExtras), c_i32p]),
    "igmk_slot_sizes": (C.c_int, [_vp, C.c_int32, c_i32p, c_i64p, c_i64p]),
    "igmk_upload": (C.c_int, [_vp]),
    "igmk_reset_coordinates": (C.c_int, [_vp]),
    "igmk_run": (C.c_int, [_vp, C.POINTER(Protocol)]),
    "igmk_single_point": (C.c_int, [_vp, C.c_double, C.c_double, C.c_double]),
    "igmk_single_point_f32": (C.c_int, [_vp, C.c_double, C.c_double, C.c_double]),
    "igmk_md_forces": (C.c_int, [_vp, C.POINTER(Protocol), C.c_int32]),
    "igmk_fetch": (C.c_int, [_vp, C.c_int32, c_f32p, C.POINTER(Result)]),
    "igmk_fetch_forces": (C.c_int, [_vp, C.c_int32, c_f64p]),
    "igmk_layout_stats": (C.c_int, [_vp, C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_double)]),
    "igmk_fetch_population": (C.c_int, [_vp, C.c_int32, C.c_int32, C.c_int32, c_f32p, C.POINTER(Result)]),
    "igmk_violation_ratios": (C.c_int, [_vp, C.c_int32, c_f64p, c_f64p]),
    "igmk_last_timings": (C.c_int, [_vp, c_f64p, c_f64p, c_i64p]),
    "igmk_single_point_noalloc": (C.c_int, [C.c_int, _vp, C.c_int32, _vp, _vp, _vp, _vp, _vp, _vp, C.c_int32, c_f64p,
                                            C.c_double, C.c_float, C.c_double, C.c_double, C.c_double, _vp, _vp, _vp]),
    "igmk_pop_create": (C.c_int, [C.c_int, C.c_int32, C.c_int32, C.c_int32, c_f32p, c_i32p, c_i32p, c_i32p,
                                  C.POINTER(_vp)]),
    "igmk_pop_destroy": (C.c_int, [_vp]),
    "igmk_pop_upload": (C.c_int, [_vp, c_f32p]),
    "igmk_pop_download": (C.c_int, [_vp, c_f32p]),
    "igmk_pop_coords_devptr": (C.c_int, [_vp, C.POINTER(_vp)]),
    "igmk_pop_actdist": (C.c_int, [_vp, C.c_int64, c_i32p, c_i32p, c_f64p, c_f64p, C.c_double,
                                   C.c_int64, c_i32p, c_i32p, c_f32p, c_f32p, c_i64p]),
    "igmk_pop_assign": (C.c_int, [_vp, C.c_int64, c_i32p, c_i32p, c_f32p, c_u32p]),
    "igmk_pop_attach_coords": (C.c_int, [_vp, _vp]),
    "igmk_pop_last_kernel_ms": (C.c_double, [_vp]),
    "igmk_pop_set_pairs": (C.c_int, [_vp, C.c_int64, c_i32p, c_i32p, c_f64p, c_f64p]),
    "igmk_pop_actdist_launch": (C.c_int, [_vp, C.c_double]),
    "igmk_pop_actdist_collect": (C.c_int, [_vp, C.c_int64, c_i32p, c_i32p, c_f32p, c_f32p, c_i64p]),
    "igmk_pop_shard_count": (C.c_int, [_vp, C.c_double, _vp]),
    "igmk_pop_shard_rank": (C.c_int, [_vp, C.c_int32, _vp, _vp, _vp]),
    "igmk_pop_shard_hist": (C.c_int, [_vp, C.c_int32, _vp, _vp]),
    "igmk_pop_shard_pick": (C.c_int, [_vp, C.c_int32, _vp, _vp, _vp]),
    "igmk_pop_contact_freq": (C.c_int, [_vp, C.c_int64, c_i32p, c_i32p, C.c_double, c_f32p]),
    "igmk_pop_damid_actdist": (C.c_int, [_vp, C.c_int64, c_i32p, c_f64p, c_f64p, C.c_double, C.c_int32, c_f64p,
                                         C.c_int64, c_i32p, c_f32p, c_f32p, c_i64p]),
    "igmk_pop_damid_assign": (C.c_int, [_vp, C.c_int64, c_i32p, c_f32p, C.c_double, c_f64p, c_u32p]),
    "igmk_pop_fish_ranks": (C.c_int, [_vp, C.c_int64, c_i32p, c_i32p, c_f32p, c_f32p, c_i32p, c_i32p]),
    "igmk_pop_sprite_rg2": (C.c_int, [_vp, C.c_int32, c_i32p, c_i32p, c_i32p, c_f32p, C.POINTER(C.c_int8), c_i32p]),
    "igmk_pop_sprite_select": (C.c_int, [_vp, C.c_int32, c_i32p, c_i32p, c_i32p, c_i32p, c_i32p, C.c_int32, c_f32p, c_i32p, c_i32p]),
    "igmk_pop_set_rows": (C.c_int, [_vp, C.c_int64, c_i32p, c_i32p, c_f32p]),
    "igmk_pop_assign_launch": (C.c_int, [_vp]),
    "igmk_pop_assign_collect": (C.c_int, [_vp, c_u32p]),
    "igmk_pop_set_row_bounds": (C.c_int, [_vp, C.c_int64, c_f64p, C.c_double]),
    "igmk_submit_population_dev": (C.c_int, [_vp, C.POINTER(System), _vp, C.c_int32, C.c_int32, c_i64p, c_i32p]),
    "igmk_debug_fetch_array": (C.c_int, [_vp, C.c_int32, _vp, C.c_int64, c_i64p]),
}


def declared_symbols():
    """Every entry point include/igmk.h declares (kept in sync by a test)."""
    return sorted(_SIGNATURES)


def lib_path():
    return _LIB_PATH


def load():
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.isfile(_LIB_PATH):
        raise RuntimeError(
            "igm_b200: CUDA library %s not found. Build it with "
            "`python -c 'import __graft_entry__ as g; g.build()'` (nvcc, sm_100a). "
            "There is no CPU fallback." % _LIB_PATH)
    lib = C.CDLL(_LIB_PATH)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    _LIB = lib
    return lib


def check(status, what=""):
    if status != 0:
        msg = load().igmk_last_error()
        raise RuntimeError("igm_b200 %s failed (status %d): %s"
                           % (what, status, msg.decode("utf8", "replace") if msg else "?"))
