"""The bond layout the MD kernel walks (md2_build_owner in igmk.cu: oriented bonds, inbox slots, sliced-ELL blocks A / B),
checked on the CPU through a staging-only handle against the bond list that was submitted: what k_md2 relies on without
testing it at run time."""
import numpy as np

from igm_b200 import synth
from igm_b200.kernel import MStepBatch, STAGING_ONLY
from igm_b200._abi import PackedSystem, BOND_UPPER, BOND_LOWER

T = 768                     # threads of k_md2 (MD2S_T, md2.cuh)
SYM, TSHIFT, NONE = 0x20000, 18, 0xffff


def _systems():
    g = synth.Genome([40, 30, 25, 20], n_autosomes=3)
    R = 2200.0
    n = len(g)
    rad = synth.bead_radii(n, R, 0.3)
    rng = np.random.default_rng(5)
    out = []
    for s in range(3):
        base = synth.pack_structure(synth.territories(g, R, rng), rad, g, nucleus_radius=R, seed=s + 1, struct_id=s, n_extra_dummies=1)
        hi = rng.integers(0, n, 900)
        hj = (hi + rng.integers(1, n - 1, 900)) % n
        hub = 11 + s                                          # more incoming bonds than an inbox holds
        others = np.setdiff1d(np.arange(n), [hub])[:60]
        bi = np.concatenate([base.bond_i, hi, np.full(60, hub), [n + 1] * 4, [7]])
        bj = np.concatenate([base.bond_j, hj, others, [3, 9, hub, 20], [7]])
        r0 = np.concatenate([base.bond_r0, np.full(900, 4.0 * float(rad[0])), np.full(60, 3.0 * float(rad[0])), [900.0] * 4, [10.0]])
        st = np.concatenate([base.bond_style, np.full(900, BOND_UPPER), np.full(60, BOND_LOWER if s == 1 else BOND_UPPER),
                             [BOND_UPPER, BOND_LOWER, BOND_UPPER, BOND_UPPER], [BOND_UPPER]])
        out.append(PackedSystem(base.xyz, base.radius, base.kind, bi, bj, np.ones(bi.size), r0, st.astype(np.uint8),
                                base.env_abck, base.env_off, base.env_members, 1.0, seed=s + 1, struct_id=s))
    return out


def test_oriented_bond_layout_covers_every_bond_exactly_once():
    systems = _systems()
    with MStepBatch(STAGING_ONLY) as b:
        for ps in systems:
            b.submit(ps)
        perm = b.debug_array(10, np.uint16)
        eblk = b.debug_array(11, np.int32).reshape(-1, 4)
        items = b.debug_array(12, np.uint32)
        mdesc = b.debug_array(9, np.uint8)
        bond_ptr = b.debug_array(5, np.int32)
        btype = b.debug_array(7, np.float32).reshape(-1, 2)
        desc_raw = b.debug_array(0, np.uint8)
    ns = len(systems)
    K = perm.size // (ns * T)
    assert perm.size == ns * K * T and eblk.shape[0] == ns * K * (T // 32)
    md = np.frombuffer(mdesc.tobytes(), dtype=np.dtype([("ell", "<i8"), ("perm", "<i4"), ("desc", "<i4"), ("K", "<i4"), ("pad", "<i4")]))
    sd = len(desc_raw) // ns
    ptr_off = 0
    for s, ps in enumerate(systems):
        n = ps.xyz.shape[0]
        pm = perm[md["perm"][s]:md["perm"][s] + K * T]
        atoms = pm[pm != NONE]
        assert sorted(atoms.tolist()) == list(range(n))                        # a permutation of the atoms
        pos = np.full(n, -1); pos[pm[pm != NONE]] = np.nonzero(pm != NONE)[0]
        btype_off = int(np.frombuffer(desc_raw[s * sd:(s + 1) * sd].tobytes()[24:28], dtype="<i4")[0])
        type0_upper = btype[btype_off, 1] > 0
        # decode every item of every warp-slot block
        got = {}                                  # (owner atom, partner atom, type) -> list of (sym, q, block)
        inbox = {}                                # receiver atom -> slots used
        eb = eblk[md["desc"][s]:md["desc"][s] + K * (T // 32)]
        base = int(md["ell"][s])
        for gi, (offa, nba, offb, nbb) in enumerate(eb):
            for blk, (off, nb) in enumerate(((offa, nba), (offb, nbb))):
                blkitems = items[base + off: base + off + nb * 128].reshape(nb, 32, 4)
                for lane in range(32):
                    p = gi * 32 + lane
                    a = pm[p]
                    for v in blkitems[:, lane, :].ravel():
                        q, partner_pos, sym, t = int(v) & 15, (int(v) >> 4) & 0x1fff, bool(int(v) & SYM), int(v) >> TSHIFT
                        if partner_pos == p:                      # padding: an item to oneself (a real self bond is never listed)
                            assert t == 0 and (q == 15 or sym)
                            continue
                        assert a != NONE
                        o = int(pm[partner_pos])
                        assert o != NONE
                        if blk == 0:
                            assert type0_upper and t == 0 and not sym
                        got.setdefault((int(a), o, t), []).append((sym, q))
                        if not sym:
                            assert q < 15
                            assert q not in inbox.setdefault(o, set()), "two bonds share an inbox slot"
                            inbox[o].add(q)
        # against the submitted bonds
        frozen = ps.kind == 1                                  # KIND_STATIC
        want = {}
        # type ids: distinct (r0, k, style) ordered by use count, ties by first appearance (igmk_submit)
        keys, first, use = [], {}, {}
        for k, key in enumerate(zip(ps.bond_r0.tolist(), ps.bond_k.tolist(), ps.bond_style.tolist())):
            use[key] = use.get(key, 0) + 1
            first.setdefault(key, k)
        order = sorted(use, key=lambda kk: (-use[kk], first[kk]))
        tid = {kk: r for r, kk in enumerate(order)}
        for i, j, key in zip(ps.bond_i.tolist(), ps.bond_j.tolist(), zip(ps.bond_r0.tolist(), ps.bond_k.tolist(), ps.bond_style.tolist())):
            if i == j:
                continue
            want[(min(i, j), max(i, j), tid[key])] = want.get((min(i, j), max(i, j), tid[key]), 0) + 1
        seen = {}
        for (a, o, t), lst in got.items():
            for sym, q in lst:
                key = (min(a, o), max(a, o), t)
                rec = seen.setdefault(key, {"oriented": 0, "sym": {}})
                if sym:
                    rec["sym"][a] = rec["sym"].get(a, 0) + 1
                else:
                    rec["oriented"] += 1
                    assert not frozen[a] and not frozen[o]
                assert not frozen[a]                                   # a frozen particle lists nothing
        assert set(seen) == set(want)
        for key, cnt in want.items():
            rec = seen[key]
            a, o, _ = key
            if frozen[a] or frozen[o]:
                moving = [x for x in (a, o) if not frozen[x]]
                assert rec["oriented"] == 0 and all(rec["sym"].get(x, 0) == cnt for x in moving)
            else:
                n_sym = rec["sym"].get(a, 0)
                assert n_sym == rec["sym"].get(o, 0)                   # a SYM bond is listed at both ends
                assert rec["oriented"] + n_sym == cnt                  # every bond exactly once
        assert all(len(v) <= 15 for v in inbox.values())
        assert max(len(v) for v in inbox.values()) == 15              # the hub's inbox is full, the rest of its bonds are SYM
        ptr_off += n + 1
    assert bond_ptr.size == ptr_off
