"""igm.model.Model -> flat arrays for the C ABI.

Walks a Model exactly the way the reference's LAMMPS front end does
(``LammpsModel.from_model``, igm/model/kernel/lammps_model.py:253-301): particles by
``ptype``, forces by ``ftype``.  Only duck typing is used (``.particles[i].pos/.r/.ptype``,
``.forces[k].ftype/.i/.j/.d/.k/.semiaxes/.particle_ids``) so the unmodified igm classes and
light stand-ins in the tests both work.

Differences from LammpsModel that do not change results: a DUMMY_STATIC particle stays
ONE atom (LAMMPS needs clones because of its 20-bonds-per-atom limit,
lammps_model.py:186,303-312; forces on frozen atoms are discarded either way), so the
particle index IS the atom index and no ``imap`` is needed.
"""
from itertools import compress
from operator import attrgetter

import numpy as np

from ._abi import PackedSystem, KIND_BEAD, KIND_STATIC, KIND_CENTROID, BOND_UPPER, BOND_LOWER, IGMK_MAX_ENVELOPES

# igm/model/forces.py:12-17, igm/model/particle.py:10-12
EXCLUDED_VOLUME, HARMONIC_UPPER_BOUND, HARMONIC_LOWER_BOUND, ENVELOPE, GENERAL_ENVELOPE, NUCL_EXCLUDED_VOLUME = range(6)
NORMAL, DUMMY_STATIC, DUMMY_DYNAMIC = range(3)

_KIND_OF_PTYPE = {NORMAL: KIND_BEAD, DUMMY_STATIC: KIND_STATIC, DUMMY_DYNAMIC: KIND_CENTROID}
_FTYPE, _BOND_I, _BOND_J, _BOND_K, _BOND_D = (attrgetter(a) for a in ("ftype", "i", "j", "k", "d"))


def script_seed(opts, model_id, step_no):
    """The seed create_lammps_script derives (lammps.py:159-165).  When ``seed`` is
    missing the reference draws a random one; so do we."""
    seed = opts.get("seed", None)
    if seed is None:
        seed = np.random.randint(1, 9007991)
    return ((int(seed) * int(model_id) * int(step_no)) % 9190037) + 1


def pack_model(model, seed=1):
    """-> PackedSystem for ``model`` (an igm.model.Model or a duck-typed equivalent).

    A 2 Mb structure carries ~30 000 Force objects; attribute access is batched (``np.fromiter`` over an
    ``attrgetter`` per field) so that packing costs ~15 ms instead of ~50 ms per structure."""
    particles = model.particles
    n = len(particles)
    xyz = np.array([p.pos for p in particles], dtype=np.float32).reshape(n, 3)
    rad = np.array([p.r for p in particles], dtype=np.float32)
    try:
        kind = np.array([_KIND_OF_PTYPE[int(p.ptype)] for p in particles], dtype=np.uint8)
    except KeyError:
        raise ValueError("Unknown particle type")                # lammps_model.py:276

    forces = model.forces
    ftypes = np.fromiter(map(_FTYPE, forces), dtype=np.int64, count=len(forces))
    env_abck, env_members, env_off = [], [], [0]
    evfactor = -1.0                                             # no Steric restraint applied
    if (ftypes == GENERAL_ENVELOPE).any():
        raise NotImplementedError(
            "igm_b200: voxel-map envelopes (ExpEnvelope / fix volumetricrestraint, "
            "igm/model/kernel/lammps.py:305-310) are not implemented by the b200 kernel")
    for k in np.nonzero(ftypes == ENVELOPE)[0]:
        f = forces[k]
        a, b, c = (float(v) for v in f.semiaxes)
        env_abck.append((a, b, c, float(f.k)))
        env_members.extend(int(v) for v in f.particle_ids)
        env_off.append(len(env_members))
    for k in np.nonzero(ftypes == EXCLUDED_VOLUME)[0]:
        evfactor = float(forces[k].k)                           # lammps_model.py:282-283 (the last one wins)
    # any other ftype is silently ignored by LammpsModel.from_model as well (:285-301)
    is_bond = (ftypes == HARMONIC_UPPER_BOUND) | (ftypes == HARMONIC_LOWER_BOUND)
    nb = int(is_bond.sum())
    if nb:
        bonds = list(compress(forces, is_bond))                 # model order kept
        bi = np.fromiter(map(_BOND_I, bonds), dtype=np.int32, count=nb)
        bj = np.fromiter(map(_BOND_J, bonds), dtype=np.int32, count=nb)
        bk = np.fromiter(map(_BOND_K, bonds), dtype=np.float64, count=nb)
        br0 = np.fromiter(map(_BOND_D, bonds), dtype=np.float64, count=nb)
        bs = np.where(ftypes[is_bond] == HARMONIC_UPPER_BOUND, BOND_UPPER, BOND_LOWER).astype(np.uint8)
    else:
        bi, bj, bk, br0, bs = [], [], [], [], []
    if len(env_abck) > IGMK_MAX_ENVELOPES:
        raise ValueError("igm_b200: at most %d envelope forces per model" % IGMK_MAX_ENVELOPES)
    return PackedSystem(xyz, rad, kind, bi, bj, bk, br0, bs,
                        np.array(env_abck, dtype=np.float64).reshape(-1, 4), env_off, env_members,
                        evfactor, seed=seed, struct_id=int(model.id))
