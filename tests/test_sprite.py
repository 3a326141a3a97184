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
