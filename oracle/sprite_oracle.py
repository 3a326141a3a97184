"""CPU restatement of the SPRITE A-step numerics.  TEST INFRASTRUCTURE ONLY.

``get_rgs2`` follows igm/cython_compiled/cpp_sprite_assignment.cpp:49-143 operation by operation
in float32.  ``reference_get_rgs2`` loads the reference's OWN native code, compiled by
oracle/build_ref.sh into oracle/_ref/ (travels to the GPU box), when it is there.  Pinned by the
known answers of igm/cython_compiled/tests.py:9,21 and tests/golden/sprite_golden.npz."""
import glob
import importlib.util
import os

import numpy as np

F = np.float32


def _rg2(pts):
    n = len(pts)
    m = np.zeros(3, dtype=F)
    for p in pts:
        m = (m + p).astype(F)
    m = (m / F(n)).astype(F)
    rg = F(0)
    for p in pts:
        d = (p - m).astype(F)
        sq = (d * d).astype(F)
        rg = F(rg + F(F(sq[0] + sq[1]) + sq[2]))
    return F(rg / F(n))


def get_rgs2(crds, copies_num):
    """crds [B][N][3] float32 (alternatives of region 0 first), copies_num [M] -> rg2s, best, copy_idxs [N][M]"""
    crds = np.asarray(crds, dtype=F)
    n_struct = crds.shape[1]
    M = len(copies_num)
    ncomb = int(np.prod(copies_num))
    start = np.concatenate([[0], np.cumsum(copies_num)])
    rg2s = np.zeros(n_struct, dtype=F)
    idx = np.zeros((n_struct, M), dtype=np.int32)
    best, best_v = -1, F(100000000)
    for s in range(n_struct):
        mv, mi = F(100000000), [-1] * M
        for k in range(ncomb):
            kk, sel = k, []
            for i in range(M):
                sel.append(kk % copies_num[i]); kk //= copies_num[i]
            v = _rg2([crds[start[i] + sel[i], s] for i in range(M)])
            if v < mv:
                mv, mi = v, sel
        rg2s[s] = mv
        idx[s] = mi
        if mv < best_v:
            best_v, best = mv, s
    return rg2s, best, idx


def reference_get_rgs2():
    """The reference's compiled get_rgs2, or None when oracle/_ref has not been built."""
    so = glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "sprite*.so"))
    if not so:
        return None
    spec = importlib.util.spec_from_file_location("sprite", so[0])
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.get_rgs2


def reduce_assignment(rg2_values, structure_indexes, n_struct, batch_size, kT, seed):
    """SpriteAssignmentStep.reduce (igm/steps/SpriteAssignmentStep.py:176-272) statement by statement on in-memory batch
    results, drawing from the GLOBAL numpy state like the reference (seeded here so that it can be replayed).
    -> (assignment, pos)"""
    np.random.seed(seed)
    n_clusters = len(rg2_values)
    n_batches = n_clusters // batch_size + (1 if n_clusters % batch_size != 0 else 0)
    random_order = np.random.permutation(range(n_batches))
    occupancy = np.zeros(n_struct, dtype=np.int32)
    assignment = np.zeros(n_clusters, dtype=np.int32)
    positions = np.zeros(n_clusters, dtype=np.int32)
    aveN = float(n_clusters) / n_struct
    stdN = np.sqrt(aveN)
    for batch_id in random_order:
        lo = batch_id * batch_size
        results = zip(rg2_values[lo:lo + batch_size], structure_indexes[lo:lo + batch_size])
        for i, (best_rg2s, curr_idx) in enumerate(results):
            ci = i + batch_id * batch_size
            if best_rg2s[0] < 0:
                pos = 0
                si = -1
            else:
                best_rgs = np.sqrt(best_rg2s)
                structure_penalizations = np.clip(occupancy[curr_idx] - aveN, 0., None) / stdN
                E = (best_rgs - best_rgs[0]) / kT + structure_penalizations
                P = np.cumsum(np.exp(-(E - E[0])))
                Z = P[-1]
                e = np.random.rand() * Z
                pos = np.searchsorted(P, e, side='left')
                si = curr_idx[pos]
                occupancy[si] += 1
            assignment[ci] = si
            positions[ci] = pos
    return assignment, positions


def reference_module():
    """the reference's compiled Cython module (get_rgs2, compute_gyration_radius), or None"""
    so = glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "sprite*.so"))
    if not so:
        return None
    spec = importlib.util.spec_from_file_location("sprite", so[0])
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod
