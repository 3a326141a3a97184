"""north_star criterion 3: population restraint-satisfaction statistics after a full A/M iteration on
the GPU against the same iteration with the CPU oracle doing the M-step (same assigned restraints,
same protocol, independent arithmetic and random streams): imposed counts identical, violated
fractions within 1 % of the imposed restraints, and the per-structure statistics of the two optimised populations
(log energies, violated restraints, radius of gyration over 64 structures) within 4.5 standard errors of their means."""
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
    S = 64
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
    systems, gpu_xyz, cpu_xyz = [], [], []
    for s in range(S):
        m = mask[:, s]
        ps = synth.pack_structure(crd[:, s, :], rad, g, row[m], col[m], r0[m], nucleus_radius=R, seed=int(seeds[s]), struct_id=s)
        x, res = ff.optimize(ps, prot)
        ratios = vo.bond_ratios(x.astype(np.float32), ps.bond_i, ps.bond_j, ps.bond_r0, ps.bond_k, ps.bond_style)
        n_imp_cpu += ratios.size; n_vio_cpu += int(np.count_nonzero(ratios > TOL))
        systems.append(ps); cpu_xyz.append(x.astype(np.float32))
        xg = ps.xyz.copy(); xg[:len(g)] = out["coordinates"][:, s, :]
        gpu_xyz.append(xg)
        assert np.isfinite(out["coordinates"][:, s, :]).all()
    assert n_imp_gpu == n_imp_cpu and n_imp_gpu > S * (len(g) - 7)
    assert abs(n_vio_gpu - n_vio_cpu) <= STAT_TOL * n_imp_cpu, (n_vio_gpu, n_vio_cpu, n_imp_cpu)
    assert n_vio_gpu <= 0.02 * n_imp_gpu                      # both end (nearly) violation free
    # per-structure energies, violated restraints and size of the two populations: means within 4.5 standard errors
    from helpers import population_stats, assert_population_stats_match
    assert_population_stats_match(population_stats(systems, gpu_xyz), population_stats(systems, cpu_xyz),
                                  floors={"n_violated": 0.1, "log_e_total": 0.01, "log_e_bond": 0.01, "e_pair_per_bead": 1e-4, "rg": 0.5})
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


def test_file_to_file_iteration(tmp_path):
    """population .hss + matrix .hcs -> actdist.hdf5 + new .hss (ActivationDistanceStep + ModelingStep on disk):
    the actdist file holds exactly the oracle's rows, a second iteration picks up its probabilities as plast,
    the population file carries the new coordinates, nstruct/nbead and the violation summary."""
    import json
    from igm_b200 import files, h5lite
    from igm_b200.iteration import hic_iteration_files
    g = synth.Genome([16, 12, 9, 6], n_autosomes=3)
    R = 1500.0
    rad = synth.bead_radii(len(g), R, 0.3)
    S = 12
    crd = synth.population_territories(g, S, R=R, seed=41)
    pi, pj, pw = synth.hic_pairs(g, 0.004, seed=6)
    # the matrix as a .hcs: CSR of the upper triangle over haploid bins (pairs arrive in row-major order)
    indptr = np.zeros(g.n_hap + 1, dtype=np.int32)
    np.add.at(indptr, pi + 1, 1)
    hcs_tree = {"matrix": {"data": pw.astype(np.float32), "indices": pj.astype(np.int32), "indptr": np.cumsum(indptr).astype(np.int32),
                           "diagonal": np.zeros(g.n_hap, dtype=np.float32)},
                "index": {"chrom": g.chrom[:g.n_hap], "copy": np.zeros(g.n_hap, dtype=np.int32),
                          "copy_index": json.dumps({str(h): [h] for h in range(g.n_hap)})}}
    hss_tree = {"coordinates": crd, "radii": rad, "summary": "{}",
                "index": {"chrom": g.chrom, "copy": g.copy, "copy_index": json.dumps({str(h): [int(b) for b in v] for h, v in g.copy_index.items()})},
                "envelope": {"shape": "sphere", "params": np.int64(R)},
                h5lite.ATTRS: {"nbead": np.int64(len(g)), "nstruct": np.int64(S), "violation": np.float64(0), "version": np.int32(2)}}
    hss_in, hcs, hss_out, ad1, ad2 = (str(tmp_path / n) for n in ("in.hss", "m.hcs", "out.hss", "actdist1.hdf5", "actdist2.hdf5"))
    h5lite.write(hcs, hcs_tree)
    h5lite.write(hss_in, hss_tree)
    opts = synth.demo_protocol_options()
    opts["custom_annealing_protocol"]["mdsteps"] = [200, 300, 300, 200]
    opts["custom_annealing_protocol"]["relax"]["mdsteps"] = 30
    sigma = float(np.sort(pw)[pw.size // 2])
    out = hic_iteration_files(hss_in, hcs, hss_out, ad1, opts, sigma, sigma, seeds=np.arange(S, dtype=np.int64) + 1)
    keep = pw.astype(np.float32) >= np.float32(sigma)
    want = ao.actdist_rows(pi[keep], pj[keep], pw[keep].astype(np.float32).astype(np.float64), np.zeros(int(keep.sum())),
                           crd, rad, g.chrom, g.copy_index, 2.0)
    got = files.read_actdist(ad1)
    assert all(np.array_equal(a.view(np.uint32), b.view(np.uint32)) for a, b in zip(got, want)) and got[0].size > 20
    new = files.read_hss(hss_out)
    assert new["nstruct"] == S and new["nbead"] == len(g) and np.array_equal(new["coordinates"], out["coordinates"])
    assert np.isfinite(new["coordinates"]).all() and not np.array_equal(new["coordinates"], crd)
    summary = json.loads(new["tree"]["summary"])
    mask = ao.assign(crd, *got[:3])
    n_poly = synth.polymer_bonds(g, rad)[0].size
    assert summary["n_imposed"] == S * n_poly + int(mask.sum()) + S * len(g)         # polymer + Hi-C + one envelope term per bead
    assert summary["byrestraint"]["Polymer+HiC"]["n_imposed_hic"] == int(mask.sum())
    assert new["tree"][h5lite.ATTRS]["violation"] == summary["n_violations"] / summary["n_imposed"]
    # second iteration at a lower sigma: plast comes from the first actdist file (rows between first copies)
    sigma2 = float(np.sort(pw)[pw.size // 4])
    hic_iteration_files(hss_out, hcs, str(tmp_path / "out2.hss"), ad2, opts, sigma2, sigma2, last_actdist=ad1,
                        seeds=np.arange(S, dtype=np.int64) + 1)
    keep2 = pw.astype(np.float32) >= np.float32(sigma2)
    pl = files.hic_pairs_from_matrix(files.read_hcs(hcs), sigma2, sigma2, got)[3]
    assert pl.size == int(keep2.sum()) and np.count_nonzero(pl) > 5
    want2 = ao.actdist_rows(pi[keep2], pj[keep2], pw[keep2].astype(np.float32).astype(np.float64), pl,
                            new["coordinates"], rad, g.chrom, g.copy_index, 2.0)
    got2 = files.read_actdist(ad2)
    assert all(np.array_equal(a.view(np.uint32), b.view(np.uint32)) for a, b in zip(got2, want2))
