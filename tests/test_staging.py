"""Host-side staging of the M-step batch without a GPU (IGMK_STAGING_ONLY handles): what pack +
create_lammps_data do before any LAMMPS process exists (lammps.py:429-431).  No compute happens here --
every call that needs the device must fail loudly on such a handle."""
import numpy as np
import pytest

from igm_b200 import synth, _abi
from igm_b200.kernel import MStepBatch, STAGING_ONLY
from oracle import astep_oracle as ao


def _population(S=70, seed=12):
    g = synth.Genome([14, 11, 8, 6], n_autosomes=3)
    R = 1400.0
    rad = synth.bead_radii(len(g), R, 0.3)
    crd = synth.population_territories(g, S, R=R, seed=seed)
    rng = np.random.default_rng(seed)
    row = rng.integers(0, len(g) - 1, size=400).astype(np.int32)
    col = (row + 1 + rng.integers(0, 12, size=400)).clip(max=len(g) - 1).astype(np.int32)
    keep = row != col
    row, col = row[keep], col[keep]
    dist = rng.uniform(300.0, 1500.0, size=row.size).astype(np.float32)        # activation distances: some structures inside
    mask = ao.assign(crd, row, col, dist)
    bits = ao.pack_bits(mask)
    r0 = (2.0 * (rad[row] + rad[col])).astype(np.float64)
    tmpl = synth.pack_structure(np.zeros((len(g), 3), dtype=np.float32), rad, g, nucleus_radius=R, seed=0, struct_id=100)
    return g, rad, R, crd, row, col, r0, mask, bits, tmpl


def test_submit_population_stages_what_per_structure_submits_stage():
    g, rad, R, crd, row, col, r0, mask, bits, tmpl = _population()
    S = crd.shape[1]
    with MStepBatch(STAGING_ONLY) as a, MStepBatch(STAGING_ONLY) as b:
        assert a.submit_population(tmpl, crd, len(g), row, col, r0, 1.0, bits) == 0 and len(a) == S
        for s in range(S):
            m = mask[:, s]
            b.submit(synth.pack_structure(crd[:, s, :], rad, g, row[m], col[m], r0[m], nucleus_radius=R, seed=s, struct_id=100 + s))
        assert [a.slot_sizes(s) for s in range(S)] == [b.slot_sizes(s) for s in range(S)]
        assert sum(a.slot_sizes(s)[1] for s in range(S)) == S * tmpl.bond_i.size + int(mask.sum()) and mask.sum() > 50
        ka, items_a, ell_a, _ = a.layout_stats()
        kb, items_b, ell_b, _ = b.layout_stats()
        assert (ka, items_a, ell_a) == (kb, items_b, ell_b) and ka == 1
        assert items_a == 2 * (S * tmpl.bond_i.size + int(mask.sum()))          # every bond from both ends
        assert items_a <= ell_a <= 1.6 * items_a + 128 * 32 * S                 # sliced-ELL padding stays bounded
        # chunks with a word offset into the same bit matrix stage the same structures (what run_population does)
        with MStepBatch(STAGING_ONLY) as c:
            assert c.submit_population(tmpl, crd, len(g), row, col, r0, 1.0, bits, s_begin=32, n_struct=38, bits_word0=1) == 0
            assert [c.slot_sizes(k) for k in range(38)] == [a.slot_sizes(32 + k) for k in range(38)]
        # a second call appends
        assert a.submit_population(tmpl, crd, len(g), row, col, r0, 1.0, bits, s_begin=64, n_struct=6, bits_word0=2) == S
        assert a.slot_sizes(S + 5) == a.slot_sizes(69)


def test_staging_only_handle_refuses_device_work():
    g, rad, R, crd, row, col, r0, mask, bits, tmpl = _population(S=3)
    prot = _abi.make_protocol(synth.demo_protocol_options())
    with MStepBatch(STAGING_ONLY) as a:
        a.submit_population(tmpl, crd, len(g), row, col, r0, 1.0, bits)
        for call in (a.upload, lambda: a.run(prot), a.single_point, lambda: a.violation_ratios(0)):
            with pytest.raises(RuntimeError, match="STAGING_ONLY"):
                call()
        with pytest.raises(RuntimeError):
            a.fetch(0)                                   # nothing has been run
        a.clear()
        assert len(a) == 0 or a.slot_sizes  # handle still usable


def test_submit_population_rejects_bad_rows():
    g, rad, R, crd, row, col, r0, mask, bits, tmpl = _population(S=5)
    bad = row.copy(); bad[len(bad) // 2] = len(g) + 7
    with MStepBatch(STAGING_ONLY) as a:
        with pytest.raises(RuntimeError, match="out of range"):
            a.submit_population(tmpl, crd, len(g), bad, col, r0, 1.0, bits)
        with pytest.raises(ValueError):
            a.submit_population(tmpl, crd, len(g), row, col, r0, 1.0, bits, s_begin=0, n_struct=5, bits_word0=1)
