"""SPRITE A-step numerics: restatement and CUDA kernel against the reference's native get_rgs2
(its known answers, igm/cython_compiled/tests.py:9,21; a golden fixture; and, where oracle/_ref
holds the compiled reference, the reference itself on fresh inputs)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import sprite_oracle as so

C1 = np.array([[[1, 0, 0]], [[-1, 0, 0]], [[0, -1, 0]], [[0, 1, 0]]], dtype=np.float32)
C3 = np.array([[[1, 0, 0], [0.1, 0, 0]], [[0.5, 0, 0], [1, 0, 0]], [[0, -0.5, 0], [-1, 0, 0]],
               [[0, 0.5, 0], [-0.1, 0, 0]]], dtype=np.float32)


@pytest.fixture(scope="module")
def sg():
    return np.load(os.path.join(GOLDEN, "sprite_golden.npz"))


def _clusters(sg):
    cl, rg, bd = sg["cl_off"], sg["rg_off"], sg["beads"]
    return [[list(bd[rg[i]:rg[i + 1]]) for i in range(cl[c], cl[c + 1])] for c in range(len(cl) - 1)]


def test_known_answers_oracle():
    r1 = so.get_rgs2(C1, [1, 1, 1, 1])
    assert np.allclose(r1[0], [1.0]) and r1[1] == 0 and np.array_equal(r1[2], [[0, 0, 0, 0]])
    r3 = so.get_rgs2(C3, [2, 2])
    assert np.allclose(r3[0], [0.125, 0.01]) and r3[1] == 1 and np.array_equal(r3[2], [[1, 0], [0, 1]])


def test_oracle_matches_golden_bit_for_bit(astep_golden, sg):
    crd = astep_golden["coordinates"]
    for c, regs in enumerate(_clusters(sg)[:12]):
        flat = [b for r in regs for b in r]
        rg2s, best, idx = so.get_rgs2(crd[flat], [len(r) for r in regs])
        assert np.array_equal(rg2s.view(np.uint32), sg["rg2_%d" % c].view(np.uint32))
        assert best == int(sg["best_%d" % c]) and np.array_equal(idx, sg["idx_%d" % c])


def _pop(g):
    from igm_b200.astep import Population
    p = Population(g["coordinates"].shape[1], g["radii"], g["chrom"], g["copy_off"], g["copy_beads"])
    p.upload(np.ascontiguousarray(g["coordinates"]))
    return p


@pytest.mark.gpu
def test_gpu_matches_golden(astep_golden, sg):
    with _pop(astep_golden) as pop:
        out = pop.sprite_rg2(_clusters(sg))
    for c, (rg2s, best, idx) in enumerate(out):
        assert np.array_equal(rg2s.view(np.uint32), sg["rg2_%d" % c].view(np.uint32)), c
        assert best == int(sg["best_%d" % c]) and np.array_equal(idx, sg["idx_%d" % c])


@pytest.mark.gpu
def test_gpu_known_answers_and_live_reference():
    from igm_b200.astep import Population
    # tests.py:9 -- four single-copy beads, one structure
    with Population(1, np.ones(4, dtype=np.float32), np.zeros(4, dtype=np.int32), np.arange(5, dtype=np.int32),
                    np.arange(4, dtype=np.int32)) as pop:
        pop.upload(C1)
        rg2s, best, idx = pop.sprite_rg2([[[0], [1], [2], [3]]])[0]
        assert rg2s[0] == 1.0 and best == 0 and np.array_equal(idx, [[0, 0, 0, 0]])
    # tests.py:21 -- two beads with two alternatives each, two structures
    with Population(2, np.ones(4, dtype=np.float32), np.zeros(4, dtype=np.int32), np.arange(5, dtype=np.int32),
                    np.arange(4, dtype=np.int32)) as pop:
        pop.upload(C3)
        rg2s, best, idx = pop.sprite_rg2([[[0, 1], [2, 3]]])[0]
        assert np.allclose(rg2s, [0.125, 0.01]) and best == 1 and np.array_equal(idx, [[1, 0], [0, 1]])
    ref = so.reference_get_rgs2()
    if ref is None:
        pytest.skip("oracle/_ref not built: the reference's native code is not available here")
    rng = np.random.default_rng(5)
    crd = (rng.normal(size=(40, 300, 3)) * 900).astype(np.float32)
    clusters = []
    for c in range(25):
        regs, used = [], 0
        for _ in range(int(rng.integers(1, 7))):
            k = int(rng.integers(1, 3))
            regs.append(list(rng.choice(40, k, replace=False)))
        clusters.append(regs)
    with Population(300, np.ones(40, dtype=np.float32), np.zeros(40, dtype=np.int32), np.arange(41, dtype=np.int32),
                    np.arange(40, dtype=np.int32)) as pop:
        pop.upload(crd)
        out = pop.sprite_rg2(clusters)
    for regs, (rg2s, best, idx) in zip(clusters, out):
        flat = [b for r in regs for b in r]
        w = ref(np.ascontiguousarray(crd[flat]), np.array([len(r) for r in regs], dtype=np.int32))
        assert np.array_equal(rg2s.view(np.uint32), w[0].view(np.uint32)) and best == w[1] and np.array_equal(idx, w[2])


def _mini_clusters(g, n, seed):
    """clusters of haploid segment ids: sizes 2 + Geom(0.3), drawn inside windows of one to four chromosomes"""
    rng = np.random.default_rng(seed)
    n_hap = len(g["copy_off"]) - 1
    chrom = g["chrom"][:n_hap]
    out = []
    for _ in range(n):
        size = 2 + rng.geometric(0.3)
        chs = rng.choice(np.unique(chrom), rng.integers(1, 5), replace=False)
        pool = np.nonzero(np.isin(chrom, chs))[0]
        out.append(rng.choice(pool, min(size, pool.size), replace=False).astype(np.int64))
    return out


@pytest.mark.gpu
def test_sprite_select_equals_the_reference_compute_gyration_radius(astep_golden):
    """Population.sprite_select = compute_gyration_radius (sprite.pyx:104-283, the reference's own compiled module from
    oracle/_ref, its np.random.choice replayed through the seed) + the keep-best selection of
    SpriteAssignmentStep.task (:148-160): Rg^2 values bit-identical, same structures, same selected beads"""
    from igm_b200 import sprite
    ref = so.reference_module()
    if ref is None:
        pytest.skip("oracle/_ref not built")
    g = astep_golden
    crd = np.ascontiguousarray(g["coordinates"])
    S = crd.shape[1]
    n_hap = len(g["copy_off"]) - 1
    copy_index = {h: [int(b) for b in g["copy_beads"][g["copy_off"][h]:g["copy_off"][h + 1]]] for h in range(n_hap)}
    index = type("Index", (), {"chrom": g["chrom"]})()
    clusters = _mini_clusters(g, 60, seed=11)
    keep = 20
    seed = 4
    batch = sprite.ClusterBatch(clusters, g["chrom"][:n_hap], g["copy_off"], g["copy_beads"], np.random.RandomState(seed), max_chrom=3)
    assert batch.skipped.any() and not batch.skipped.all()
    with _pop(g) as pop:
        best_rg2, best_struct, selected = pop.sprite_select(batch, keep_best=keep)
    np.random.seed(seed)                                          # the reference draws np.random.choice cluster by cluster, in order
    for c, cluster in enumerate(clusters):
        if batch.skipped[c]:
            assert (best_rg2[c] == -1).all() and (best_struct[c] == -1).all() and (selected[c] == -1).all()
            continue
        rg2s, _, sel = ref.compute_gyration_radius(crd, cluster, index, copy_index)
        rg2s = np.asarray(rg2s, dtype=np.float32)
        ind = np.lexsort((np.arange(S), rg2s))[:keep]             # argpartition + argsort (:148-149), ties by structure index
        assert np.array_equal(best_rg2[c].view(np.uint32), rg2s[ind].view(np.uint32)), c
        assert np.array_equal(best_struct[c], ind)
        assert np.array_equal(selected[c], np.asarray(sel)[ind])


@pytest.mark.gpu
def test_sprite_restraints_through_the_mstep(astep_golden):
    """clusters -> selection -> Gibbs assignment -> centroids + bonds as per-structure atoms of the population submit: the
    batch holds exactly the systems a per-structure build gives (sizes, double-precision single point vs the oracle) and
    the MD kernel moves the centroids (DUMMY_DYNAMIC)"""
    from igm_b200 import sprite, synth, restraints_pop as rp
    from igm_b200._abi import PackedSystem, KIND_CENTROID, BOND_UPPER
    from igm_b200.kernel import MStepBatch
    from oracle import ff
    from helpers import short_protocol
    g = astep_golden
    crd = np.ascontiguousarray(g["coordinates"][:, :48, :])
    S, n = crd.shape[1], crd.shape[0]
    n_hap = len(g["copy_off"]) - 1
    idx = type("G", (), {"chrom": g["chrom"], "copy": g["copy"], "copy_off": g["copy_off"], "copy_beads": g["copy_beads"],
                         "__len__": lambda self: n})()
    rad = g["radii"]
    rs = np.random.RandomState(2)
    clusters = _mini_clusters(g, 90, seed=5)
    batch = sprite.ClusterBatch(clusters, g["chrom"][:n_hap], g["copy_off"], g["copy_beads"], rs)
    with _pop({**{k: g[k] for k in ("radii", "chrom", "copy_off", "copy_beads")}, "coordinates": crd}) as pop:
        best_rg2, best_struct, selected = pop.sprite_select(batch, keep_best=10)
    assignment, pos = sprite.assign_clusters(best_rg2, best_struct, S, rs, batch_size=10, kT=100.0)
    beads = [selected[c][pos[c]] for c in range(len(clusters))]
    tmpl = synth.pack_structure(np.zeros((n, 3), dtype=np.float32), rad, idx, nucleus_radius=5500.0, seed=0, struct_id=0)
    atoms, bonds = sprite.centroid_restraints(crd, rad, assignment, beads, 0, tmpl.n_atoms, 0.05, k=0.8)
    assert atoms[0][-1] == (assignment >= 0).sum() and np.diff(atoms[0]).max() >= 2
    extras = rp.extras(S, bonds=bonds, atoms=atoms)
    none_i = np.zeros(0, dtype=np.int32)
    prot = short_protocol(md=60, relax=10)
    with MStepBatch(0) as b:
        b.submit_population(tmpl, crd, n, none_i, none_i, np.zeros(0), 1.0, np.zeros((0, 1), dtype=np.uint32), extras=extras,
                            seeds=np.arange(S, dtype=np.int64) + 3)
        b.single_point()
        systems = []
        for s in range(S):
            nc = int(atoms[0][s + 1] - atoms[0][s])
            sl = slice(bonds[0][s], bonds[0][s + 1])
            xyz = np.vstack([crd[:, s, :], np.zeros((1, 3), np.float32), atoms[1][atoms[0][s]:atoms[0][s + 1]]])
            ps = PackedSystem(xyz, np.concatenate([tmpl.radius, np.zeros(nc, np.float32)]),
                              np.concatenate([tmpl.kind, np.full(nc, KIND_CENTROID, np.uint8)]),
                              np.concatenate([tmpl.bond_i, bonds[1][sl]]), np.concatenate([tmpl.bond_j, bonds[2][sl]]),
                              np.concatenate([tmpl.bond_k, bonds[3][sl]]), np.concatenate([tmpl.bond_r0, bonds[4][sl]]),
                              np.concatenate([tmpl.bond_style, bonds[5][sl]]), tmpl.env_abck, tmpl.env_off, tmpl.env_members,
                              tmpl.evfactor, seed=3 + s, struct_id=s)
            systems.append(ps)
            assert b.slot_sizes(s) == (ps.n_atoms, ps.bond_i.size, ps.env_members.size)
            f_ref, e_ref = ff.single_point(ps)
            _, res = b.fetch(s)
            assert res.final_energy == pytest.approx(e_ref["total"], rel=1e-10)
            assert np.abs(b.fetch_forces(s) - f_ref).max() <= 1e-10 * np.abs(f_ref).max()
        b.run(prot)
        moved = 0
        for s in range(S):
            xyz, res = b.fetch(s)
            assert res.status == 0 and np.isfinite(xyz).all() and xyz.shape[0] == systems[s].n_atoms
            moved += int(np.any(xyz[n + 1:] != systems[s].xyz[n + 1:]))
        assert moved >= S // 2
