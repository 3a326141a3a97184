"""north_star criterion 3: population restraint-satisfaction statistics after a full A/M iteration on
the GPU against the same iteration with the CPU oracle doing the M-step (same assigned restraints,
same protocol, independent arithmetic and random streams): imposed counts identical, violated
fractions within 1 % of the imposed restraints (tolerance stated here), energies of the same order."""
import numpy as np
import pytest

from igm_b200 import synth
from igm_b200.iteration import hic_iteration
from oracle import ff, astep_oracle as ao, viol_oracle as vo

pytestmark = pytest.mark.gpu

TOL = 0.05            # optimization/violation_tolerance of the demo config
STAT_TOL = 0.01       # allowed difference of the violated fraction between the two implementations


def test_am_iteration_statistics_match_cpu():
    g = synth.Genome([16, 12, 9, 6], n_autosomes=3)
    R = 1500.0
    rad = synth.bead_radii(len(g), R, 0.3)
    S = 24
    crd = synth.population_territories(g, S, R=R, seed=21)
    pi, pj, pw = synth.hic_pairs(g, 0.004, seed=4)
    pl = np.zeros(pi.size)
    opts = synth.demo_protocol_options()
    opts["custom_annealing_protocol"]["mdsteps"] = [300, 600, 600, 400]
    opts["custom_annealing_protocol"]["relax"]["mdsteps"] = 50
    template = synth.pack_structure(np.zeros((len(g), 3), dtype=np.float32), rad, g, nucleus_radius=R, seed=0, struct_id=0)
    seeds = np.arange(S, dtype=np.int64) + 3
    out = hic_iteration(crd, g, rad, template, (pi, pj, pw, pl), opts, seeds=seeds, violations=True, tol=TOL)
    row, col, dist, prob = out["rows"]
    # the A-step half is bit-exact
    want = ao.actdist_rows(pi, pj, pw, pl, crd, rad, g.chrom, g.copy_index, 2.0)
    assert all(np.array_equal(a.view(np.uint32), w.view(np.uint32)) for a, w in zip(out["rows"], want))
    mask = ao.assign(crd, row, col, dist)
    r0 = (2.0 * (rad[row] + rad[col])).astype(np.float64)
    from igm_b200 import _abi
    prot = _abi.make_protocol(opts)
    n_imp_gpu = sum(v[0] for v in out["violations"]); n_vio_gpu = sum(v[1] for v in out["violations"])
    n_imp_cpu = n_vio_cpu = 0
    e_gpu, e_cpu = [], []
    for s in range(S):
        m = mask[:, s]
        ps = synth.pack_structure(crd[:, s, :], rad, g, row[m], col[m], r0[m], nucleus_radius=R, seed=int(seeds[s]), struct_id=s)
        x, res = ff.optimize(ps, prot)
        ratios = vo.bond_ratios(x.astype(np.float32), ps.bond_i, ps.bond_j, ps.bond_r0, ps.bond_k, ps.bond_style)
        n_imp_cpu += ratios.size; n_vio_cpu += int(np.count_nonzero(ratios > TOL))
        e_cpu.append(res.final_energy); e_gpu.append(out["infos"][s]["final-energy"])
        assert np.isfinite(out["coordinates"][:, s, :]).all()
    assert n_imp_gpu == n_imp_cpu and n_imp_gpu > S * (len(g) - 7)
    assert abs(n_vio_gpu - n_vio_cpu) <= STAT_TOL * n_imp_cpu, (n_vio_gpu, n_vio_cpu, n_imp_cpu)
    assert n_vio_gpu <= 0.02 * n_imp_gpu                      # both end (nearly) violation free
    assert np.median(e_gpu) <= 10 * np.median(e_cpu) + 1.0 and np.median(e_cpu) <= 10 * np.median(e_gpu) + 1.0
    # every bead ends inside the nucleus
    assert (np.linalg.norm(out["coordinates"], axis=2) + rad[:, None] <= R * 1.02).all()


def test_relax_init_removes_overlaps_and_keeps_the_polymer():
    """RelaxInit.task for a block of structures (RelaxInit.py:86-178): territories in, relaxed
    coordinates out -- inside the envelope, consecutive beads within the polymer bond length, the
    excluded-volume energy far below where it started; identical to the same structures submitted one
    by one (the population path only changes who assembles them)."""
    from igm_b200.iteration import relax_init
    from igm_b200.kernel import MStepBatch
    from igm_b200 import _abi
    g = synth.Genome([16, 12, 9, 6], n_autosomes=3)
    R = 1500.0
    rad = synth.bead_radii(len(g), R, 0.3)
    S = 9
    crd = synth.population_territories(g, S, R=R, seed=31)
    opts = synth.demo_protocol_options()
    opts["custom_annealing_protocol"]["mdsteps"] = [300, 600, 600, 400]
    opts["custom_annealing_protocol"]["relax"]["mdsteps"] = 50
    template = synth.pack_structure(np.zeros((len(g), 3), dtype=np.float32), rad, g, nucleus_radius=R, seed=0, struct_id=0)
    seeds = np.arange(S, dtype=np.int64) + 11
    new, infos = relax_init(crd, template, opts, seeds=seeds)
    assert new.shape == crd.shape and len(infos) == S and np.isfinite(new).all()
    assert (np.linalg.norm(new, axis=2) + rad[:, None] <= R * 1.02).all()
    with MStepBatch(0) as b:
        for s in range(S):
            b.submit(synth.pack_structure(crd[:, s, :], rad, g, nucleus_radius=R, seed=int(seeds[s]), struct_id=s))
        b.single_point()
        e0 = [b.fetch(s)[1].pair_energy for s in range(S)]
        b.run(_abi.make_protocol(opts))
        for s in range(S):
            xyz, res = b.fetch(s)
            assert np.array_equal(xyz[:len(g)].view(np.uint32), new[:, s, :].view(np.uint32))
            assert infos[s]["final-energy"] == res.final_energy
            assert res.pair_energy < 0.05 * e0[s] + 1.0
    # polymer bonds of the template hold: consecutive beads of a chain no further apart than r0 (+5 %)
    d = np.linalg.norm(new[template.bond_i] - new[template.bond_j], axis=2)
    assert (d <= template.bond_r0[:, None] * 1.05).all()
