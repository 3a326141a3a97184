#!/usr/bin/env python
"""Golden fixture for the per-structure restraint builders: the force lists the UNMODIFIED reference classes
``igm.restraints.Fish`` and ``igm.restraints.Damid`` put on a Model (imported from /root/reference through
refstub.py; run in the build container only):

    python tests/golden/make_golden_restraints.py  ->  tests/golden/restraints_golden.npz

Population = the mini genome of astep_golden.npz (100 demo structures); FISH targets = sorted samples of the
population's own distances plus noise (what an experimental distribution looks like to the code)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.abspath(os.path.join(HERE, "..", ".."))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import refstub                      # noqa: E402


def main():
    refstub.import_reference()
    from igm.model import Model, Particle
    from igm.restraints import Fish, Damid
    from igm.model.forces import HarmonicLowerBound, HarmonicUpperBound, EllipticEnvelope
    fas = sys.modules["igm.steps.FishAssignmentStep"]

    g = np.load(os.path.join(HERE, "astep_golden.npz"))
    fg = np.load(os.path.join(HERE, "fish_golden.npz"))
    dg = np.load(os.path.join(HERE, "damid_golden.npz"))
    crd, rad = g["coordinates"], g["radii"]
    off, beads = g["copy_off"], g["copy_beads"]
    copy_index = {h: [int(b) for b in beads[off[h]:off[h + 1]]] for h in range(len(off) - 1)}
    n_bead, S = crd.shape[0], crd.shape[1]
    index = refstub.FakeIndex(g["chrom"], g["copy"] if "copy" in g.files else np.zeros(n_bead), copy_index, [])
    rng = np.random.default_rng(21)
    probes = fg["probes"][:12]
    pairs = np.stack([fg["pair_i"][:15], fg["pair_j"][:15]], axis=1)
    # FishAssignmentStep.task: target[idx] with idx = rank of every structure (:220-225, :247-252)
    tgt = {}
    for name, items, rows_min, rows_max, imin, imax in (("radial", probes, fg["rmin"][:12], fg["rmax"][:12], fg["rimin"][:12], fg["rimax"][:12]),
                                                        ("pair", pairs, fg["pmin"][:15], fg["pmax"][:15], fg["pimin"][:15], fg["pimax"][:15])):
        tmin = np.sort(rows_min + rng.normal(0, 50.0, rows_min.shape), axis=1)
        tmax = np.sort(rows_max + rng.normal(0, 50.0, rows_max.shape), axis=1)
        tgt[name + "_min_sorted"], tgt[name + "_max_sorted"] = tmin, tmax
        tgt[name + "_min"] = np.take_along_axis(tmin, imin, axis=1)
        tgt[name + "_max"] = np.take_along_axis(tmax, imax, axis=1)
    refstub.FAKE_FILES["golden_fish_assignment"] = {"probes": probes, "pairs": pairs, "radial_min": tgt["radial_min"],
                                                    "radial_max": tgt["radial_max"], "pair_min": tgt["pair_min"], "pair_max": tgt["pair_max"]}
    refstub.FAKE_FILES["golden_damid_rows"] = {"loc": dg["ellipsoid_bead"], "dist": dg["ellipsoid_dist"], "prob": dg["ellipsoid_prob"]}
    structs = [0, 3, 17, 50, 99]
    tol, ck = 120.0, 2.5
    out = {"structs": np.array(structs), "tol": tol, "kspring": ck, "probes": probes, "pairs": pairs}
    out.update(tgt)
    semi, cr, kd = dg["ellipsoid_param"], float(dg["ellipsoid_cr"]), 0.35
    for s in structs:
        model = Model(uid=s)
        for b in range(n_bead):
            model.addParticle(crd[b, s], rad[b], Particle.NORMAL)
        model.addRestraint(Damid("golden_damid_rows", contact_range=cr, nuclear_radius=0, k=kd, shape="ellipsoid", semiaxes=semi))
        n_after_damid = len(model.forces)
        model.addRestraint(Fish("golden_fish_assignment", index, s, "rRpP", tolerance=tol, kspring=ck))
        f = model.forces[0]
        assert isinstance(f, EllipticEnvelope) and n_after_damid == 1
        out["damid_members_%d" % s] = np.array(f.particle_ids, dtype=np.int32)
        out["damid_abck_%d" % s] = np.array(list(f.semiaxes) + [f.k], dtype=np.float64)
        rows = []
        for f in model.forces[1:]:
            style = 0 if isinstance(f, HarmonicUpperBound) else 1
            assert isinstance(f, (HarmonicUpperBound, HarmonicLowerBound))
            rows.append((style, f.i, f.j, f.d, f.k))
        out["fish_forces_%d" % s] = np.array(rows, dtype=np.float64)
        out["dummies_%d" % s] = np.array([k for k, p in enumerate(model.particles) if p.ptype == Particle.DUMMY_STATIC])
    # ---- Sprite._apply (igm/restraints/sprite.py:41-69): centroids + bead-centroid bonds of the clusters assigned to a structure
    from igm.restraints import Sprite
    rng = np.random.default_rng(5)
    n_cl = 40
    sizes = 2 + rng.geometric(0.3, n_cl)
    indptr = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    selected = np.concatenate([rng.choice(n_bead, int(k), replace=False) for k in sizes]).astype(np.int32)
    assignment = rng.choice(np.array(structs + [7, 8]), n_cl).astype(np.int32)
    refstub.FAKE_FILES["golden_sprite_assignment"] = {"assignment": assignment, "indptr": indptr, "selected": selected}
    out.update(sprite_assignment=assignment, sprite_indptr=indptr, sprite_selected=selected, sprite_volume_occupancy=0.05, sprite_k=1.7)
    for s in structs:
        model = Model(uid=s)
        for b in range(n_bead):
            model.addParticle(crd[b, s], rad[b], Particle.NORMAL)
        model.addRestraint(Sprite("golden_sprite_assignment", 0.05, s, k=1.7))
        out["sprite_forces_%d" % s] = np.array([(f.i, f.j, f.d, f.k) for f in model.forces], dtype=np.float64).reshape(-1, 4)
        cen = [p for p in model.particles if p.ptype == Particle.DUMMY_DYNAMIC]
        out["sprite_centroids_%d" % s] = np.array([p.pos for p in cen], dtype=np.float32).reshape(-1, 3)
        assert all(isinstance(f, HarmonicUpperBound) for f in model.forces)
    np.savez_compressed(os.path.join(HERE, "restraints_golden.npz"), **out)
    print("restraints golden: %d FISH forces per structure, DamID members %s"
          % (out["fish_forces_0"].shape[0], [out["damid_members_%d" % s].size for s in structs]))


if __name__ == "__main__":
    main()
