"""The drop-in boundary with the reference's REAL classes: an ``igm.model.Model`` built by the unmodified
``Steric / Envelope / Polymer / interHiC / intraHiC`` restraints (the pure-Python part of the reference, staged under
oracle/_ref/igm_py by oracle/build_ref.sh; h5py replaced by an in-memory stand-in), ``Model.optimize`` patched with the
three lines of INTEGRATION.md section 2, run through ``B200Controller`` -- against the same structures packed directly.
And the A-step binding: ``steps_adapter.actdist_task`` with the file protocol of ``ActivationDistanceStep.task``."""
import os
import sys

import numpy as np
import pytest

from conftest import ROOT, GOLDEN
from igm_b200 import synth, files, steps_adapter
from igm_b200.controller import B200Controller
from igm_b200.kernel import MStepBatch, optimize, kernel_opts
from igm_b200.pack import script_seed
from igm_b200 import _abi
from oracle import astep_oracle as ao

pytestmark = pytest.mark.gpu
STAGED = os.path.join(ROOT, "oracle", "_ref", "igm_py")


def _short_options():
    opts = synth.demo_protocol_options()
    cap = opts["custom_annealing_protocol"]
    cap["mdsteps"] = [120] * 4
    cap["relax"]["mdsteps"] = 20
    opts["seed"] = 77
    return opts


@pytest.mark.skipif(not os.path.isfile(os.path.join(STAGED, "igm", "model", "model.py")), reason="oracle/_ref/igm_py not staged")
def test_real_model_and_restraints_through_the_b200_kernel(astep_golden):
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import refstub
    refstub.import_staged(STAGED)
    from igm.model import Model, Particle
    from igm.restraints import Steric, Envelope, Polymer, interHiC, intraHiC

    def b200_optimize(self, cfg):                     # the branch added to Model.optimize
        if cfg['optimization']["kernel"] == "b200":
            from igm_b200 import kernel as b200
            return b200.optimize(self, cfg)
    Model.optimize = b200_optimize

    g = astep_golden
    crd, rad = np.ascontiguousarray(g["coordinates"]), g["radii"]
    n = crd.shape[0]
    index = refstub.FakeIndex(g["chrom"], g["copy"], {}, [])
    refstub.FAKE_FILES["real_actdist"] = {"row": g["row"], "col": g["col"], "dist": g["dist"], "prob": g["prob"]}
    R = 5500.0
    cfg = {"optimization": {"kernel": "b200", "optimizer_options": _short_options(), "kernel_opts": {"b200": {"device": 0}}},
           "runtime": {"step_no": 3}}
    structs = [0, 3, 17, 50]

    def task(s):                                      # the part of ModelingStep.task around model.optimize (:184-264, :335)
        model = Model(uid=s)
        for b in range(n):
            model.addParticle(crd[b, s], rad[b], Particle.NORMAL)
        model.addRestraint(Steric(1.0))
        model.addRestraint(Envelope("sphere", R, 1.0))
        model.addRestraint(Polymer(index, 2.0, 1.0))
        model.addRestraint(interHiC("real_actdist", g["chrom"], 2.0, 1.0))
        model.addRestraint(intraHiC("real_actdist", g["chrom"], 2.0, 1.0))
        info = model.optimize(cfg)
        return model, info

    ctl = B200Controller()
    out = ctl.map(task, structs)
    assert ctl.last_batcher.batches == [len(structs)]                 # one launch for all parked optimize calls
    # the same structures packed directly from the arrays, same seeds (lammps.py:159-165 through script_seed)
    mask = ao.assign(crd, g["row"], g["col"], g["dist"])
    run_opts = dict(cfg["optimization"]["optimizer_options"]); run_opts.update(kernel_opts(cfg))
    prot = _abi.make_protocol(run_opts)
    gen = type("G", (), {"chrom": g["chrom"], "copy": g["copy"], "__len__": lambda self: n})()
    with MStepBatch(0) as b:
        slots = []
        for s in structs:
            inter = g["chrom"][g["row"]] != g["chrom"][g["col"]]
            m = mask[:, s]
            order = np.concatenate([np.nonzero(m & inter)[0], np.nonzero(m & ~inter)[0]])     # interHiC first, then intraHiC
            r0 = (np.float32(2.0) * (rad[g["row"][order]] + rad[g["col"][order]])).astype(np.float64)
            ps = synth.pack_structure(crd[:, s, :], rad, gen, g["row"][order], g["col"][order], r0, nucleus_radius=R,
                                      seed=script_seed(run_opts, s, 3 + 2), struct_id=s)
            slots.append(b.submit(ps))
        b.run(prot)
        for (model, info), slot, s in zip(out, slots, structs):
            xyz, res = b.fetch(slot)
            got = np.array([p.pos for p in model.particles], dtype=np.float32)
            assert len(model.particles) == n + 1 and len(model.forces) == b.slot_sizes(slot)[1] + 2   # + Steric + Envelope
            assert np.array_equal(got.view(np.uint32), xyz.view(np.uint32)), s
            assert info["final-energy"] == res.final_energy and info["thermo"]["E_bond"] == res.bond_energy
            assert set(info) == {"final-energy", "pair-energy", "bond-energy", "thermo", "md-time"}
            assert not np.array_equal(got[:n], crd[:, s, :])


def test_actdist_task_binding_follows_the_reference_file_protocol(astep_golden, tmp_path):
    """<batch>.in.npy -> actdist_task -> <batch>.out.tmp, parsed the way ActivationDistanceStep.reduce does (:238-249)"""
    g = astep_golden
    crd = np.ascontiguousarray(g["coordinates"])
    hss = str(tmp_path / "pop.hss")
    ci = {str(h): [int(v) for v in g["copy_beads"][g["copy_off"][h]:g["copy_off"][h + 1]]] for h in range(len(g["copy_off"]) - 1)}
    tree = {"radii": g["radii"], "index": {"chrom": g["chrom"], "copy": g["copy"], "copy_index": __import__("json").dumps(ci)}}
    files.write_hss(hss, tree, coordinates=crd)
    cfg = {"optimization": {"structure_output": hss}, "restraints": {"Hi-C": {"contact_range": 2.0}}}
    n = g["pair_i"].size
    cuts = [0, n // 3, n // 3, n]                                       # three batches, the middle one empty
    for b in range(3):
        sl = slice(cuts[b], cuts[b + 1])
        batch = [(int(i), int(j), pw, pl) for i, j, pw, pl in zip(g["pair_i"][sl], g["pair_j"][sl], g["pwish"][sl], g["plast"][sl])]
        np.save(str(tmp_path / ("%d.in.npy" % b)), batch)
    for b in range(3):
        assert steps_adapter.actdist_task(b, cfg, str(tmp_path)) is None
    shape = [("row", "int32"), ("col", "int32"), ("dist", "float32"), ("prob", "float32")]        # actdist_shape, :33-38
    parts = []
    for b in (0, 2):
        part = np.genfromtxt(str(tmp_path / ("%d.out.tmp" % b)), dtype=shape)
        parts.append(np.array([part], dtype=shape) if part.ndim == 0 else part)
    assert os.path.getsize(str(tmp_path / "1.out.tmp")) == 0
    got = np.concatenate(parts)
    for name in ("row", "col", "dist", "prob"):
        assert np.array_equal(got[name].view(np.uint32), g[name].view(np.uint32)), name
    steps_adapter.release()
