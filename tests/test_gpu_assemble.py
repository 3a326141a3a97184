"""the assignment -> M-step hand-off on the device (igmk_submit_population_dev, assemble.cuh)
builds, from the population's slab, rows and bit matrix, the batch igmk_submit_population + igmk_upload build on
the host from the downloaded bit matrix -- array by array, byte by byte -- and therefore the same optimised
structures, violation reports and A/M iterations."""
import numpy as np
import pytest

from igm_b200 import synth, _abi
from igm_b200.astep import Population
from igm_b200.kernel import MStepBatch, run_population
from igm_b200._abi import PackedSystem, BOND_UPPER, BOND_LOWER

pytestmark = pytest.mark.gpu

ARRAYS = ["desc", "xyzr", "xyzr0", "kind", "envmask", "bond_ptr", "bond_items", "btype_f", "btype_d", "md2_desc", "perm",
          "eblk", "ell_items"]


def _population(g, rad, S, R, sigma, seed):
    crd = synth.population_territories(g, S, R=R, seed=seed)
    pi, pj, pw = synth.hic_pairs(g, sigma, seed=seed + 1)
    pop = Population(S, rad, g.chrom, g.copy_off, g.copy_beads)
    pop.upload(crd)
    row, col, dist, prob = pop.actdist(pi, pj, pw, np.zeros(pi.size), 2.0)
    pop.set_rows(row, col, dist)
    pop.assign_launch()
    return crd, pop, row, col


def _compare(host, dev):
    for k, name in enumerate(ARRAYS):
        a, b = host.debug_array(k), dev.debug_array(k)
        assert a.size == b.size, "%s: %d bytes on the host path, %d on the device path" % (name, a.size, b.size)
        if not np.array_equal(a, b):
            bad = np.nonzero(a != b)[0]
            raise AssertionError("%s differs in %d of %d bytes, first at byte %d (host %s, device %s)"
                                 % (name, bad.size, a.size, bad[0], a[bad[0]:bad[0] + 16], b[bad[0]:bad[0] + 16]))


def _template_with_everything(g, rad, R):
    """polymer bonds + a lower bound, a bond to a frozen dummy, a bond of a bead to itself and a hub bead with 40 upper
    bounds (more than an inbox holds), radii of three sizes (several bound distances among the Hi-C rows)"""
    n = len(g)
    t = synth.pack_structure(np.zeros((n, 3), dtype=np.float32), rad, g, nucleus_radius=R, seed=5, struct_id=100, n_extra_dummies=1)
    rng = np.random.default_rng(3)
    hub = 7
    others = rng.choice(np.setdiff1d(np.arange(n), [hub]), 40, replace=False)
    bi = np.concatenate([t.bond_i, [2, 11, 5], np.full(40, hub), [n + 1] * 3])
    bj = np.concatenate([t.bond_j, [9, n, 5], others, [3, 4, hub]])
    bk = np.concatenate([t.bond_k, [0.7, 1.0, 1.0], np.full(40, 1.0), [1.0] * 3])
    r0 = np.concatenate([t.bond_r0, [300.0, 800.0, 10.0], np.full(40, 2.0 * (rad[0] + rad[0])), [900.0, 900.0, 400.0]])
    st = np.concatenate([t.bond_style, [BOND_LOWER, BOND_UPPER, BOND_UPPER], np.full(40, BOND_UPPER), [BOND_UPPER, BOND_LOWER, BOND_UPPER]])
    return PackedSystem(t.xyz, t.radius, t.kind, bi, bj, bk, r0, st.astype(np.uint8), t.env_abck, t.env_off, t.env_members,
                        t.evfactor, seed=5, struct_id=100)


@pytest.mark.parametrize("s_begin,n_struct", [(0, 32), (32, 38), (5, 61)])
def test_device_assembly_is_byte_identical_small(s_begin, n_struct):
    g = synth.Genome([16, 12, 9, 6], n_autosomes=3)
    R = 1500.0
    rad = synth.bead_radii(len(g), R, 0.3) * (1.0 + 0.08 * (np.arange(len(g)) % 3)).astype(np.float32)
    S = 70
    crd, pop, row, col = _population(g, rad, S, R, 0.004, 21)
    with pop:
        tmpl = _template_with_everything(g, rad, R)
        r0 = (2.0 * (rad[row] + rad[col])).astype(np.float64)
        assert np.unique(r0).size > 3
        bits = pop.assign_collect()
        seeds = np.arange(n_struct, dtype=np.int64) * 7 + 1 if s_begin else None
        with MStepBatch() as host, MStepBatch() as dev:
            if s_begin % 32 == 0:
                host.submit_population(tmpl, crd, len(g), row, col, r0, 1.5, bits, s_begin=s_begin, n_struct=n_struct, seeds=seeds,
                                       bits_word0=s_begin // 32)
            else:       # the host path takes bit 0 of word bits_word0 for structure s_begin: re-pack the columns
                from igm_b200.astep import unpack_bits
                m = unpack_bits(bits, S)[:, s_begin:s_begin + n_struct]
                packed = np.packbits(np.pad(m, ((0, 0), (0, (-n_struct) % 32))), axis=1, bitorder="little").view(np.uint32)
                host.submit_population(tmpl, crd, len(g), row, col, r0, 1.5, packed, s_begin=s_begin, n_struct=n_struct, seeds=seeds)
            host.upload()
            pop.set_row_bounds(r0, 1.5)
            assert dev.submit_population_device(tmpl, pop, s_begin, n_struct, seeds=seeds)
            _compare(host, dev)
            # same batch => same trajectories, bit for bit; same violation report
            opts = synth.demo_protocol_options()
            opts["custom_annealing_protocol"]["mdsteps"] = [100, 150, 150, 100]
            opts["custom_annealing_protocol"]["relax"]["mdsteps"] = 20
            prot = _abi.make_protocol(opts)
            host.run(prot); dev.run(prot)
            for k in (0, n_struct // 2, n_struct - 1):
                xa, ra = host.fetch(k); xb, rb = dev.fetch(k)
                assert np.array_equal(xa.view(np.uint32), xb.view(np.uint32))
                assert ra.final_energy == rb.final_energy and ra.n_rebuilds == rb.n_rebuilds
                assert host.slot_sizes(k) == dev.slot_sizes(k)
                va, ea = host.violation_ratios(k); vb, eb = dev.violation_ratios(k)
                assert np.array_equal(va, vb) and np.array_equal(ea, eb)


def test_device_assembly_is_byte_identical_2mb():
    """the BASELINE configs[1] shape: 3008 beads + centre, three atoms per thread, tens of thousands of bonds per structure"""
    g = synth.genome_2mb()
    R = 5500.0
    rad = synth.bead_radii(len(g), R, 0.4)
    S = 40
    crd, pop, row, col = _population(g, rad, S, R, 0.01, 5)
    with pop:
        tmpl = synth.pack_structure(np.zeros((len(g), 3), dtype=np.float32), rad, g, nucleus_radius=R, seed=0, struct_id=0)
        r0 = (2.0 * (rad[row] + rad[col])).astype(np.float64)
        bits = pop.assign_collect()
        with MStepBatch() as host, MStepBatch() as dev:
            host.submit_population(tmpl, crd, len(g), row, col, r0, 1.0, bits, s_begin=0, n_struct=S)
            host.upload()
            pop.set_row_bounds(r0, 1.0)
            assert dev.submit_population_device(tmpl, pop, 0, S)
            assert host.slot_sizes(3)[1] > len(g) + 2000 and row.size > 1000000
            _compare(host, dev)


def test_run_population_from_the_device_equals_the_host_path():
    g = synth.Genome([16, 12, 9, 6], n_autosomes=3)
    R = 1500.0
    rad = synth.bead_radii(len(g), R, 0.3)
    S = 70
    crd, pop, row, col = _population(g, rad, S, R, 0.004, 33)
    with pop:
        tmpl = synth.pack_structure(np.zeros((len(g), 3), dtype=np.float32), rad, g, nucleus_radius=R, seed=0, struct_id=0)
        r0 = (2.0 * (rad[row] + rad[col])).astype(np.float64)
        opts = synth.demo_protocol_options()
        opts["custom_annealing_protocol"]["mdsteps"] = [100, 150, 150, 100]
        opts["custom_annealing_protocol"]["relax"]["mdsteps"] = 20
        prot = _abi.make_protocol(opts)
        a, ra, va, ta = run_population(tmpl, crd, len(g), row, col, r0, 1.0, pop.assign_collect(), prot, chunk=32, violations_tol=0.05)
        b, rb, vb, tb = run_population(tmpl, crd, len(g), row, col, r0, 1.0, None, prot, chunk=32, violations_tol=0.05, pop=pop)
        assert tb["assembled_on_device"] == 1 and ta["assembled_on_device"] == 0
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
        assert va == vb
        assert tb["h2d_bytes"] < ta["h2d_bytes"] / 10


def test_what_the_device_path_does_not_cover_falls_back():
    g = synth.genome_2mb()
    R = 5500.0
    rad = (synth.bead_radii(len(g), R, 0.4) * (1.0 + 0.01 * np.random.default_rng(2).random(len(g)))).astype(np.float32)   # every bound distance distinct
    S = 33
    crd, pop, row, col = _population(g, rad, S, R, 0.05, 41)
    with pop:
        tmpl = synth.pack_structure(np.zeros((len(g), 3), dtype=np.float32), rad, g, nucleus_radius=R, seed=0, struct_id=0)
        r0 = (2.0 * (rad[row] + rad[col])).astype(np.float64)
        assert np.unique(r0).size > 1024
        pop.set_row_bounds(r0, 1.0)
        with MStepBatch() as dev:
            assert dev.submit_population_device(tmpl, pop, 0, S) is False
            assert len(dev) == 0
        opts = synth.demo_protocol_options()
        opts["custom_annealing_protocol"]["mdsteps"] = [40, 40, 40, 40]
        opts["custom_annealing_protocol"]["relax"]["mdsteps"] = 10
        new, res, _, tim = run_population(tmpl, crd, len(g), row, col, r0, 1.0, None, _abi.make_protocol(opts), pop=pop)
        assert tim["assembled_on_device"] == 0 and np.isfinite(new).all() and all(r.status == 0 for r in res)
        # and a batch cannot be filled both ways
        with MStepBatch() as b:
            b.submit(tmpl)
            with pytest.raises(RuntimeError, match="empty"):
                b.submit_population_device(tmpl, pop, 0, S)
