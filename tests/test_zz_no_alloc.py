"""The allocation-free single point (igmk_single_point_noalloc) against the regular entry point.

Who needs it: a host that took one GPU arena at start-up and must not call cudaMalloc afterwards
(or cannot, because the card is full, or will not, because cudaMalloc/cudaFree synchronise the
device).  Such a caller hands in its own device buffers; the library's scratch is a fixed array
inside the library.  Here torch plays that caller: it owns the input and output tensors, the entry
point is called through the C ABI with their addresses, and the results must equal those of
igmk_single_point -- on consecutive calls with different structures, since the calls share the
scratch array and it keeps whatever the previous call left in it.  (Collected last on purpose.)"""
import numpy as np
import pytest

from igm_b200 import noalloc, synth
from igm_b200.kernel import MStepBatch
from oracle import ff as ff_oracle
from helpers import demo_structure, small_system

pytestmark = pytest.mark.gpu

RTOL = 1e-10          # the tolerance of the float64 single-point tests (tests/test_gpu_ff.py)


def _reference(ps):
    with MStepBatch(0) as b:
        slot = b.submit(ps)
        b.single_point()
        _, res = b.fetch(slot)
        return b.fetch_forces(slot), res


def _call(torch, ps):
    lay = noalloc.host_layout(ps)
    dev = {k: torch.from_numpy(lay[k]).cuda() for k in ("xyzr", "kind", "envmask", "bond_ptr", "bond_items", "btype")}
    forces = torch.full((ps.n_atoms, 3), float("nan"), dtype=torch.float64, device="cuda")
    energies = torch.full((11,), float("nan"), dtype=torch.float64, device="cuda")
    status = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    noalloc.single_point_noalloc(0, 0, ps.n_atoms, dev["xyzr"].data_ptr(), dev["kind"].data_ptr(), dev["envmask"].data_ptr(),
                                 dev["bond_ptr"].data_ptr(), dev["bond_items"].data_ptr(), dev["btype"].data_ptr(),
                                 lay["env_abck"], lay["evfactor"], lay["max_radius"],
                                 forces.data_ptr(), energies.data_ptr(), status.data_ptr())
    torch.cuda.synchronize()
    return forces.cpu().numpy(), energies.cpu().numpy(), int(status.item())


def _check(torch, ps):
    f_ref, res = _reference(ps)
    f, e, status = _call(torch, ps)
    assert status == 0
    f_orc, e_orc = ff_oracle.single_point(ps)               # and directly against the CPU oracle
    assert e[0] == pytest.approx(e_orc["total"], rel=RTOL) and np.abs(f - f_orc).max() <= RTOL * np.abs(f_orc).max()
    assert e[0] == pytest.approx(res.final_energy, rel=RTOL)
    assert e[1] == pytest.approx(res.pair_energy, rel=RTOL, abs=RTOL * abs(res.final_energy))
    assert e[2] == pytest.approx(res.bond_energy, rel=RTOL, abs=RTOL * abs(res.final_energy))
    for j in range(res.n_env):
        assert e[3 + j] == pytest.approx(res.env_energy[j], rel=RTOL, abs=RTOL * abs(res.final_energy))
    assert np.abs(f - f_ref).max() <= RTOL * np.abs(f_ref).max()


def test_noalloc_matches_regular_entry_on_consecutive_calls(ff_golden):
    torch = pytest.importorskip("torch")
    big = demo_structure(ff_golden, 0)                      # 3010 atoms, every force-field term
    other = demo_structure(ff_golden, 6, n_extra_bonds=500)
    _, _, small = small_system(seed=21)
    for ps in (big, small[0], other, big):                  # later calls see the earlier calls' scratch contents
        _check(torch, ps)


def test_noalloc_rejects_sizes_beyond_its_scratch():
    torch = pytest.importorskip("torch")
    z = torch.zeros(16, dtype=torch.float64, device="cuda")
    with pytest.raises(RuntimeError, match="at most %d" % noalloc.MAX_ATOMS):
        noalloc.single_point_noalloc(0, 0, noalloc.MAX_ATOMS + 1, *([z.data_ptr()] * 6), np.zeros((0, 4)), 1.0, 1.0,
                                     z.data_ptr(), z.data_ptr(), z.data_ptr())


def test_noalloc_reports_a_full_neighbour_table_and_keeps_the_callers_device():
    """more than 96 beads inside the list cutoff of one bead: the fixed scratch cannot hold the rows -> status 1 (the
    regular entry point widens its table instead); and the call must not leave another device current"""
    torch = pytest.importorskip("torch")
    g = synth.Genome([400], n_autosomes=0)
    rad = synth.bead_radii(len(g), 1500.0, 0.4)
    xyz = (np.random.default_rng(1).normal(size=(400, 3)) * 120.0).astype(np.float32)
    ps = synth.pack_structure(xyz, rad, g, nucleus_radius=1500.0)
    before = torch.cuda.current_device()
    _, _, status = _call(torch, ps)
    assert status == 1
    assert torch.cuda.current_device() == before
    f_ref, res = _reference(ps)                              # the allocating entry point handles the same input
    assert np.isfinite(res.final_energy)
    _, _, small = small_system(seed=22)                      # and the next call on the dirty scratch is right again
    _check(torch, small[0])
