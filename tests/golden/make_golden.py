#!/usr/bin/env python
"""Generates the committed golden fixtures in tests/golden/ by running the
UNMODIFIED reference (/root/reference, imported through refstub.py) and by cutting
the known-answer data out of the demo output the reference ships.

Run in the build container only:   python tests/golden/make_golden.py
Outputs:
  ff_golden.npz      10 full demo structures + the per-structure energies lmp_serial
                     reported for them (demo/demo_sample_outputs/igm-model.hss.T)
  astep_golden.npz   a mini genome (chr1, chr2, chr21, chr22, X, Y; 100 structures) cut
                     from the demo population, a set of haploid pairs with real pwish
                     values from demo/WTC11_HiC_2Mb.hcs, and what the reference's
                     get_actdist (+ text round trip) and inter/intraHiC._apply return
  lammps_golden.json the .data / .lam numbers the reference front end
                     (LammpsModel, create_lammps_data, create_lammps_script) writes for
                     a small model with every particle / force kind
"""
import io
import json
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.abspath(os.path.join(HERE, "..", ".."))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

from igm_b200 import h5lite     # noqa: E402
import refstub                     # noqa: E402
from oracle import astep_oracle    # noqa: E402

DEMO = "/root/reference/demo"


def main():
    refstub.import_reference()
    ads = sys.modules["igm.steps.ActivationDistanceStep"]
    from igm.model import Model, Particle
    from igm.restraints import Polymer, Envelope, Steric, interHiC, intraHiC
    from igm.model.forces import HarmonicLowerBound, HarmonicUpperBound, EllipticEnvelope

    hss = h5lite.File(os.path.join(DEMO, "demo_sample_outputs", "igm-model.hss.T"))
    crd = hss["coordinates"].read()
    radii = hss["radii"].read()
    chrom = hss["index/chrom"].read()
    copy = hss["index/copy"].read()
    copy_index = {int(k): v for k, v in json.loads(hss["index/copy_index"].read()).items()}
    summ = json.loads(hss["summary"].read())["bystructure"]

    # ---------------- force-field known answers -------------------------------------
    sel = [0, 1, 5, 6, 7, 37, 80, 38, 9, 69]      # incl. the largest envelope and pair energies
    np.savez_compressed(
        os.path.join(HERE, "ff_golden.npz"),
        struct_ids=np.array(sel, dtype=np.int32),
        coordinates=np.ascontiguousarray(crd[:, sel, :]),
        radii=radii, chrom=chrom, copy=copy,
        pair_energies=np.array(summ["pair_energies"])[sel],
        bond_energies=np.array(summ["bond_energies"])[sel],
        total_energies=np.array(summ["total_energies"])[sel],
        f_envelope0=np.array(summ["thermo"]["f_envelope0"])[sel],
        nucleus_radius=np.float64(5500.0),
    )

    # ---------------- mini genome for the A-step ---------------------------------------
    keep_chrom = [0, 1, 20, 21, 22, 23]
    nhap_full = len(copy_index)
    hap_keep = np.array([h for h in range(nhap_full) if chrom[h] in keep_chrom])
    bead_keep = np.array([b for b in range(len(chrom)) if chrom[b] in keep_chrom])
    bead_new = -np.ones(len(chrom), dtype=np.int64)
    bead_new[bead_keep] = np.arange(len(bead_keep))
    hap_new = {int(h): k for k, h in enumerate(hap_keep)}
    m_crd = np.ascontiguousarray(crd[bead_keep])
    m_rad = radii[bead_keep]
    m_chrom = chrom[bead_keep]
    m_copy = copy[bead_keep]
    m_copy_index = {hap_new[int(h)]: [int(bead_new[b]) for b in copy_index[int(h)]] for h in hap_keep}
    # haploid bins come first in the bead order, so chrom[bead] == chrom[hap] for bead < nhap
    assert all(m_copy_index[k][0] == k for k in m_copy_index)
    nhap = len(hap_keep)

    hcs = h5lite.File(os.path.join(DEMO, "WTC11_HiC_2Mb.hcs"))
    data = hcs["matrix/data"].read()
    indices = hcs["matrix/indices"].read()
    indptr = hcs["matrix/indptr"].read()
    rng = np.random.default_rng(20261106)
    cand = []
    for h in hap_keep:
        for q in range(indptr[h], indptr[h + 1]):
            j = int(indices[q])
            if j in hap_new and j != h and data[q] >= 0.01:
                cand.append((hap_new[int(h)], hap_new[j], data[q]))
    cand = np.array(cand, dtype=np.float64)
    intra = m_chrom[cand[:, 0].astype(int)] == m_chrom[cand[:, 1].astype(int)]
    pick = np.concatenate([rng.choice(np.nonzero(intra)[0], 450, replace=False),
                           rng.choice(np.nonzero(~intra)[0], 350, replace=False)])
    pick.sort()
    pairs = cand[pick]
    pi = pairs[:, 0].astype(np.int32)
    pj = pairs[:, 1].astype(np.int32)
    pwish = pairs[:, 2]                                # float32 values widened (the .in.npy batches)
    plast = np.where(rng.random(len(pi)) < 0.5, 0.0,
                     np.float32(np.round(rng.random(len(pi)) * pwish, 4)).astype(np.float64))
    plast[::97] = 1.0                                  # exercise the pexist >= 1 branch

    index = refstub.FakeIndex(m_chrom, m_copy, m_copy_index, [])
    fake = refstub.FakeHss(m_crd, m_rad, index)
    lines = []
    for i, j, pw, pl in zip(pi, pj, pwish, plast):
        for r in ads.get_actdist(int(i), int(j), pw, pl, fake, contactRange=2.0):
            lines.append(ads.actdist_fmt_str % r)
    parsed = np.genfromtxt(io.StringIO("\n".join(lines)), dtype=ads.actdist_shape)
    row, col, dist, prob = parsed["row"], parsed["col"], parsed["dist"], parsed["prob"]

    # the numpy restatement must agree with the reference bit for bit
    o_row, o_col, o_dist, o_prob = astep_oracle.actdist_rows(pi, pj, pwish, plast, m_crd, m_rad, m_chrom,
                                                             m_copy_index, 2.0)
    assert np.array_equal(o_row, row) and np.array_equal(o_col, col)
    assert np.array_equal(o_dist.view(np.uint32), dist.view(np.uint32))
    assert np.array_equal(o_prob.view(np.uint32), prob.view(np.uint32))

    # assignment through the reference restraint classes, 5 structures
    refstub.FAKE_FILES["golden_actdist"] = {"row": row, "col": col, "dist": dist, "prob": prob}
    assign_structs = [0, 3, 17, 50, 99]
    bonds = []
    for s in assign_structs:
        model = Model(uid=s)
        for b in range(len(m_rad)):
            model.addParticle(m_crd[b, s], m_rad[b], Particle.NORMAL)
        model.addRestraint(interHiC("golden_actdist", m_chrom, 2.0, 1.0))
        model.addRestraint(intraHiC("golden_actdist", m_chrom, 2.0, 1.0))
        got = sorted((int(f.i), int(f.j)) for f in model.forces)
        bonds.append(np.array(got, dtype=np.int32).reshape(-1, 2))
        mask = astep_oracle.assign(m_crd[:, [s], :], row, col, dist)[:, 0]
        mine = sorted((int(a_), int(b_)) for a_, b_ in zip(row[mask], col[mask]))
        assert mine == got, "assignment restatement differs from the reference (struct %d)" % s
    np.savez_compressed(
        os.path.join(HERE, "astep_golden.npz"),
        coordinates=m_crd, radii=m_rad, chrom=m_chrom, copy=m_copy,
        copy_off=np.cumsum([0] + [len(m_copy_index[k]) for k in range(nhap)]).astype(np.int32),
        copy_beads=np.concatenate([m_copy_index[k] for k in range(nhap)]).astype(np.int32),
        pair_i=pi, pair_j=pj, pwish=pwish, plast=plast,
        row=row, col=col, dist=dist, prob=prob,
        assign_structs=np.array(assign_structs, dtype=np.int32),
        **{"bonds_%d" % s: b for s, b in zip(assign_structs, bonds)}
    )
    print("astep golden: %d pairs -> %d rows; bonds/struct %s" % (len(pi), len(row), [len(b) for b in bonds]))

    # ---------------- DamID (lamina) A-step through the reference's own functions ------------
    dads = sys.modules["igm.steps.DamidActivationDistanceStep"]
    from igm.restraints import Damid
    from oracle import damid_oracle
    d_loc = np.sort(rng.choice(nhap, 150, replace=False)).astype(np.int32)
    d_pw = np.round(np.clip(rng.normal(0.3, 0.25, d_loc.size), 0, 1), 3)
    d_pl = np.where(rng.random(d_loc.size) < 0.5, 0.0, np.round(rng.random(d_loc.size) * d_pw, 5))
    d_pl[::41] = 1.0
    damid = {}
    for shape, param, cr in (("sphere", 5500.0, 0.05), ("ellipsoid", [5600.0, 5400.0, 5100.0], 0.08)):
        lines = []
        for i, pw, pl in zip(d_loc, d_pw, d_pl):
            for r in dads.get_damid_actdist(int(i), pw, pl, fake, contact_range=cr, shape=shape, nucleus_param=param):
                lines.append(dads.damid_actdist_fmt_str % r)
        parsed = np.genfromtxt(io.StringIO("\n".join(lines)), dtype=dads.damid_actdist_shape)
        o_b, o_d, o_p = damid_oracle.damid_rows(d_loc, d_pw, d_pl, m_crd, m_rad, m_copy_index, cr, shape, param)
        assert np.array_equal(o_b, parsed["loc"])
        assert np.array_equal(o_d.view(np.uint32), parsed["dist"].view(np.uint32))
        assert np.array_equal(o_p.view(np.uint32), parsed["prob"].view(np.uint32))
        damid[shape + "_bead"], damid[shape + "_dist"], damid[shape + "_prob"] = parsed["loc"], parsed["dist"], parsed["prob"]
        # the restraint's predicate on a few structures (Damid._apply_envelope)
        refstub.FAKE_FILES["golden_damid_" + shape] = {"loc": parsed["loc"], "dist": parsed["dist"], "prob": parsed["prob"]}
        semi = (param,) * 3 if shape == "sphere" else tuple(param)
        mask = damid_oracle.damid_assign(m_crd, m_rad, parsed["loc"], parsed["dist"], cr, semi)
        for s in assign_structs:
            model = Model(uid=s)
            for b in range(len(m_rad)):
                model.addParticle(m_crd[b, s], m_rad[b], Particle.NORMAL)
            model.addRestraint(Damid("golden_damid_" + shape, contact_range=cr, nuclear_radius=param if shape == "sphere" else 0.0,
                                     shape=shape, semiaxes=semi, k=0.3))
            got = sorted(int(v) for v in model.forces[-1].particle_ids)
            assert got == sorted(int(v) for v in parsed["loc"][mask[:, s]]), "DamID predicate differs (struct %d)" % s
            damid["%s_members_%d" % (shape, s)] = np.array(got, dtype=np.int32)
    np.savez_compressed(os.path.join(HERE, "damid_golden.npz"), loc=d_loc, pwish=d_pw, plast=d_pl,
                        sphere_param=np.float64(5500.0), sphere_cr=np.float64(0.05),
                        ellipsoid_param=np.array([5600.0, 5400.0, 5100.0]), ellipsoid_cr=np.float64(0.08), **damid)
    print("damid golden: %d loci -> %d rows per shape; members/struct %s" %
          (d_loc.size, damid["sphere_bead"].size, [damid["sphere_members_%d" % s].size for s in assign_structs]))

    # ---------------- FISH A-step through the reference's own helpers ----------------------------
    fas = sys.modules["igm.steps.FishAssignmentStep"]
    dip = np.array([h for h in range(nhap) if len(m_copy_index[h]) == 2])
    f_i = rng.choice(dip, 40, replace=False).astype(np.int32)
    f_j = rng.choice(dip, 40, replace=False).astype(np.int32)
    f_j[f_j == f_i] = dip[0]
    f_probe = rng.choice(dip, 25, replace=False).astype(np.int32)
    fish = {k: [] for k in ("pmin", "pmax", "pimin", "pimax", "rmin", "rmax", "rimin", "rimax")}
    n_conf = m_crd.shape[1]
    for a_, b_ in zip(f_i, f_j):
        d = fas.get_pair_dists(m_copy_index[int(a_)], m_copy_index[int(b_)], len(m_rad), n_conf, m_crd)
        mn, mx, imn, imx = fas.get_min_max_and_idx(d)
        assert len(np.unique(mn)) == n_conf and len(np.unique(mx)) == n_conf      # no ties in the fixture
        fish["pmin"].append(mn); fish["pmax"].append(mx); fish["pimin"].append(imn); fish["pimax"].append(imx)
    for a_ in f_probe:
        d = fas.get_rad_dists(m_copy_index[int(a_)], len(m_rad), n_conf, m_crd)
        mn, mx, imn, imx = fas.get_min_max_and_idx(d)
        assert len(np.unique(mn)) == n_conf and len(np.unique(mx)) == n_conf
        fish["rmin"].append(mn); fish["rmax"].append(mx); fish["rimin"].append(imn); fish["rimax"].append(imx)
    np.savez_compressed(os.path.join(HERE, "fish_golden.npz"), pair_i=f_i, pair_j=f_j, probes=f_probe,
                        **{k: np.array(v) for k, v in fish.items()})
    print("fish golden: %d pairs, %d probes" % (len(f_i), len(f_probe)))

    # ---------------- violation ratios through the reference's Force classes (M10) ---------------
    from oracle import viol_oracle
    from igm_b200.kernel import envelope_ratios
    sv = 17
    vmodel = Model(uid=sv)
    nbv = 200
    for b in range(nbv):
        vmodel.addParticle(m_crd[b, sv], m_rad[b], Particle.NORMAL)
    vb_i = rng.integers(0, nbv, 300); vb_j = (vb_i + rng.integers(1, nbv - 1, 300)) % nbv
    vb_style = rng.integers(0, 2, 300)
    vb_k = rng.choice([1.0, 0.5, 2.5], 300)
    vb_r0 = np.where(rng.random(300) < 0.6,
                     (rng.choice([2.0, 4.0, 8.0], 300) * (m_rad[vb_i] + m_rad[vb_j])).astype(np.float32).astype(np.float64),
                     rng.uniform(200.0, 6000.0, 300))          # float32-valued (restraints) and float64-valued (FISH targets)
    vb_r0[::37] = 0.0
    v_ratio = []
    for a_, b_, st, kk, d in zip(vb_i, vb_j, vb_style, vb_k, vb_r0):
        dd = np.float32(d) if float(np.float32(d)) == d else np.float64(d)
        f = (HarmonicUpperBound if st == 0 else HarmonicLowerBound)((int(a_), int(b_)), dd, float(kk))
        v_ratio.append(float(f.getViolationRatio(vmodel.particles)))
    env_specs = [((5100.0, 5100.0, 5100.0), 1.0, 510.0, list(range(nbv))),
                 ((5225.0, 4750.0, 4275.0), -0.3, 4750.0, [int(v) for v in rng.choice(nbv, 60, replace=False)])]
    v_env = []
    for semi, kk, scale, members in env_specs:
        f = EllipticEnvelope(members, 0, semi, kk, scale=scale)
        v_env.append(f.getViolationRatios(vmodel.particles))
        u = viol_oracle.envelope_u(m_crd[:nbv, sv], m_rad, members, semi, kk)
        assert np.array_equal(envelope_ratios(u, kk, scale), v_env[-1])
    o_ratio = viol_oracle.bond_ratios(m_crd[:nbv, sv], vb_i, vb_j, vb_r0, vb_k, vb_style)
    assert np.array_equal(o_ratio, np.array(v_ratio)), np.abs(o_ratio - np.array(v_ratio)).max()
    np.savez_compressed(os.path.join(HERE, "viol_golden.npz"), struct=np.int32(sv), n_bead=np.int32(nbv), bond_i=vb_i, bond_j=vb_j,
                        bond_style=vb_style, bond_k=vb_k, bond_r0=vb_r0, bond_ratio=np.array(v_ratio),
                        env0_semi=np.array(env_specs[0][0]), env0_k=np.float64(1.0), env0_scale=np.float64(510.0),
                        env0_members=np.array(env_specs[0][3]), env0_ratio=v_env[0],
                        env1_semi=np.array(env_specs[1][0]), env1_k=np.float64(-0.3), env1_scale=np.float64(4750.0),
                        env1_members=np.array(env_specs[1][3]), env1_ratio=v_env[1])
    print("violation golden: %d bonds (%d violated), envelope members violated %d / %d" %
          (len(v_ratio), np.count_nonzero(v_ratio), np.count_nonzero(v_env[0]), np.count_nonzero(v_env[1])))

    # ---------------- reference LAMMPS front end on a small model ----------------------
    from igm.model.kernel import lammps as reflammps
    from igm.model.kernel.lammps_model import LammpsModel
    s = 7
    nb = 60
    model = Model(uid=s)
    rad2 = m_rad.copy()
    rad2[10:20] = np.float32(140.5)                  # two bead types
    for b in range(nb):
        model.addParticle(m_crd[b, s], rad2[b], Particle.NORMAL, chainID=0 if b < 40 else 1)
    model.addRestraint(Steric(1.0))
    model.addRestraint(Envelope("ellipsoid", (5000.0, 4000.0, 3000.0), 1.0))
    idx60 = refstub.FakeIndex([0] * 40 + [1] * 20, [0] * 60, {}, [40, 20])
    model.addRestraint(Polymer(idx60, 2.0, 1.0))
    model.addForce(HarmonicUpperBound((3, 50), 1200.0, 0.5))
    model.addForce(HarmonicLowerBound((4, 30), 2500.0, 2.0))
    centroid = model.addParticle(np.mean(m_crd[:5, s], axis=0), 0.0, Particle.DUMMY_DYNAMIC)
    for b in range(5):
        model.addForce(HarmonicUpperBound((b, centroid), 300.0, 1.0))
    dummy = model.addParticle([0., 0., 0.], 0., Particle.DUMMY_STATIC)
    model.addForce(EllipticEnvelope([5, 6, 7], dummy, (4500.0, 3600.0, 2700.0), -0.3, scale=50.0))
    for b in range(25):                              # > MAX_BONDS: forces a dummy clone
        model.addForce(HarmonicLowerBound((b, dummy), 1000.0 + b, 1.0))
    tmp = tempfile.mkdtemp()
    opts = json.load(open(os.path.join(DEMO, "config_file.json")))["optimization"]
    run_opts = dict(opts["optimizer_options"])
    run_opts.update(opts["kernel_opts"]["lammps"])
    run_opts.update({"out": "traj", "data": os.path.join(tmp, "x.data"), "lmp": os.path.join(tmp, "x.lam"),
                     "step_no": 5})
    lm = LammpsModel(model)
    reflammps.create_lammps_data(lm, run_opts)
    reflammps.create_lammps_script(lm, run_opts)
    data_txt = open(run_opts["data"]).read()
    lam_txt = open(run_opts["lmp"]).read()

    def section(name):
        body = data_txt.split("\n%s\n\n" % name)[1].split("\n\n")[0]
        return [ln.split() for ln in body.strip().split("\n")]

    atoms = section("Atoms")
    lam_lines = lam_txt.split("\n")
    golden = {
        "particles": {"pos": [[float(v) for v in p.pos] for p in model.particles],
                      "r": [float(p.r) for p in model.particles],
                      "ptype": [int(p.ptype) for p in model.particles]},
        "forces": [],
        "model_id": s, "seed": run_opts["seed"], "step_no": 5,
        "imap": [int(v) for v in lm.imap],
        "atoms": [[int(a[0]), int(a[1]), int(a[2]), float(a[3]), float(a[4]), float(a[5])] for a in atoms],
        "bond_coeffs": [[int(a[0]), a[1], float(a[2]), float(a[3])] for a in section("Bond Coeffs")],
        "bonds": [[int(v) for v in a] for a in section("Bonds")],
        "pair_coeffs": [[int(a[0]), int(a[1]), float(a[2]), float(a[3])] for a in section("PairIJ Coeffs")],
        "user": [[int(a[0]), int(a[1]), float(a[2])] for a in section("User")],
        "script_velocity_lines": [ln for ln in lam_lines if ln.startswith("velocity")],
        "script_envelope_lines": [ln for ln in lam_lines if ln.startswith("fix envelope")],
        "script_run_lines": [ln for ln in lam_lines if ln.startswith("run ") or ln.startswith("minimize")],
        "script_rescale_lines": [ln for ln in lam_lines if "temp/rescale" in ln],
        "script_limit_lines": [ln for ln in lam_lines if "nve/limit" in ln],
        "script_group_lines": [ln for ln in lam_lines if ln.startswith("group ")],
        "script_misc_lines": [ln for ln in lam_lines if ln.split(" ")[0] in
                              ("timestep", "pair_style", "neighbor", "special_bonds", "variable")],
        "optimizer_options": opts["optimizer_options"],
    }
    for f in model.forces:
        if f.ftype == f.EXCLUDED_VOLUME:
            golden["forces"].append({"ftype": int(f.ftype), "k": float(f.k)})
        elif f.ftype == f.ENVELOPE:
            golden["forces"].append({"ftype": int(f.ftype), "k": float(f.k),
                                     "semiaxes": [float(v) for v in f.semiaxes],
                                     "particle_ids": [int(v) for v in f.particle_ids]})
        else:
            golden["forces"].append({"ftype": int(f.ftype), "i": int(f.i), "j": int(f.j),
                                     "d": float(f.d), "k": float(f.k)})
    json.dump(golden, open(os.path.join(HERE, "lammps_golden.json"), "w"), indent=0)
    print("lammps golden: %d atoms, %d bond types, %d bonds" %
          (len(golden["atoms"]), len(golden["bond_coeffs"]), len(golden["bonds"])))

    # SPRITE known answers (igm/cython_compiled/tests.py:9,21) through the reference's native code
    from igm.cython_compiled.sprite import get_rgs2
    c1 = np.array([[[1, 0, 0]], [[-1, 0, 0]], [[0, -1, 0]], [[0, 1, 0]]], dtype=np.float32)
    r1 = get_rgs2(c1, np.array([1, 1, 1, 1], dtype=np.int32))
    c3 = np.array([[[1, 0, 0], [0.1, 0, 0]], [[0.5, 0, 0], [1, 0, 0]], [[0, -0.5, 0], [-1, 0, 0]],
                   [[0, 0.5, 0], [-0.1, 0, 0]]], dtype=np.float32)
    r3 = get_rgs2(c3, np.array([2, 2], dtype=np.int32))
    assert np.allclose(r1[0], [1.0]) and r1[1] == 0
    assert np.allclose(r3[0], [0.125, 0.01]) and r3[1] == 1 and np.array_equal(r3[2], [[1, 0], [0, 1]])
    print("sprite known answers reproduce")
    # random clusters on the mini genome through the reference's native code
    rng2 = np.random.default_rng(13)
    chroms_dip = [c for c in keep_chrom if c < 22]
    sp = {}
    n_cl = 30
    cl_regions = []
    for c in range(n_cl):
        nreg = int(rng2.integers(1, 5))
        regs = []
        for _ in range(nreg):
            h = int(rng2.choice(dip if rng2.random() < 0.8 else np.arange(nhap)))
            regs.append(m_copy_index[h])
        cl_regions.append(regs)
        flat = [b for r in regs for b in r]
        crds = np.ascontiguousarray(m_crd[flat])
        rg2s, best, idx = get_rgs2(crds, np.array([len(r) for r in regs], dtype=np.int32))
        sp["rg2_%d" % c], sp["best_%d" % c], sp["idx_%d" % c] = rg2s, np.int32(best), idx
    sp["cl_off"] = np.cumsum([0] + [len(r) for r in cl_regions]).astype(np.int32)
    sp["rg_off"] = np.cumsum([0] + [len(r) for regs in cl_regions for r in regs]).astype(np.int32)
    sp["beads"] = np.array([b for regs in cl_regions for r in regs for b in r], dtype=np.int32)
    np.savez_compressed(os.path.join(HERE, "sprite_golden.npz"), **sp)
    print("sprite golden: %d clusters" % n_cl)


if __name__ == "__main__":
    main()
