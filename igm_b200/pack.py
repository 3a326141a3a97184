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
import numpy as np

from ._abi import PackedSystem, KIND_BEAD, KIND_STATIC, KIND_CENTROID, BOND_UPPER, BOND_LOWER, IGMK_MAX_ENVELOPES

# igm/model/forces.py:12-17, igm/model/particle.py:10-12
EXCLUDED_VOLUME, HARMONIC_UPPER_BOUND, HARMONIC_LOWER_BOUND, ENVELOPE, GENERAL_ENVELOPE, NUCL_EXCLUDED_VOLUME = range(6)
NORMAL, DUMMY_STATIC, DUMMY_DYNAMIC = range(3)

_KIND_OF_PTYPE = {NORMAL: KIND_BEAD, DUMMY_STATIC: KIND_STATIC, DUMMY_DYNAMIC: KIND_CENTROID}


def script_seed(opts, model_id, step_no):
    """The seed create_lammps_script derives (lammps.py:159-165).  When ``seed`` is
    missing the reference draws a random one; so do we."""
    seed = opts.get("seed", None)
    if seed is None:
        seed = np.random.randint(1, 9007991)
    return ((int(seed) * int(model_id) * int(step_no)) % 9190037) + 1


def pack_model(model, seed=1):
    """-> PackedSystem for ``model`` (an igm.model.Model or a duck-typed equivalent)."""
    n = len(model.particles)
    xyz = np.empty((n, 3), dtype=np.float32)
    rad = np.empty(n, dtype=np.float32)
    kind = np.empty(n, dtype=np.uint8)
    for i, p in enumerate(model.particles):
        xyz[i] = p.pos
        rad[i] = p.r
        try:
            kind[i] = _KIND_OF_PTYPE[int(p.ptype)]
        except KeyError:
            raise ValueError("Unknown particle type")            # lammps_model.py:276

    bi, bj, bk, br0, bs = [], [], [], [], []
    env_abck, env_members, env_off = [], [], [0]
    evfactor = -1.0                                             # no Steric restraint applied
    for f in model.forces:
        ft = int(f.ftype)
        if ft == ENVELOPE:
            a, b, c = (float(v) for v in f.semiaxes)
            env_abck.append((a, b, c, float(f.k)))
            env_members.extend(int(v) for v in f.particle_ids)
            env_off.append(len(env_members))
        elif ft == GENERAL_ENVELOPE:
            raise NotImplementedError(
                "igm_b200: voxel-map envelopes (ExpEnvelope / fix volumetricrestraint, "
                "igm/model/kernel/lammps.py:305-310) are not implemented by the b200 kernel")
        elif ft == EXCLUDED_VOLUME:
            evfactor = float(f.k)                               # lammps_model.py:282-283
        elif ft == HARMONIC_UPPER_BOUND or ft == HARMONIC_LOWER_BOUND:
            bi.append(int(f.i))
            bj.append(int(f.j))
            bk.append(float(f.k))
            br0.append(float(f.d))
            bs.append(BOND_UPPER if ft == HARMONIC_UPPER_BOUND else BOND_LOWER)
        # any other ftype is silently ignored by LammpsModel.from_model as well (:285-301)
    if len(env_abck) > IGMK_MAX_ENVELOPES:
        raise ValueError("igm_b200: at most %d envelope forces per model" % IGMK_MAX_ENVELOPES)
    return PackedSystem(xyz, rad, kind, bi, bj, bk, br0, bs,
                        np.array(env_abck, dtype=np.float64).reshape(-1, 4), env_off, env_members,
                        evfactor, seed=seed, struct_id=int(model.id))
