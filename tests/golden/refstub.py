"""Import the UNMODIFIED reference package from /root/reference in THIS container.

The reference needs h5py / alabtools / matplotlib, none of which exist here, so
tiny in-memory stand-ins are injected into ``sys.modules`` first.  Only the
golden-fixture generator (make_golden.py) uses this; nothing that runs on the
GPU box may import it (the reference tree does not exist there).
"""
import sys
import types

import numpy as np

REFERENCE_ROOT = "/root/reference"
FAKE_FILES = {}          # path -> {dataset name: ndarray}; served by the h5py stand-in


class _FakeH5File(dict):
    def __init__(self, path, mode="r", **kw):
        if path not in FAKE_FILES:
            FAKE_FILES[path] = {}
        super(_FakeH5File, self).__init__()
        self._store = FAKE_FILES[path]
        self.update(self._store)
        self.attrs = {}

    def create_dataset(self, name, data=None, **kw):
        self._store[name] = np.asarray(data)
        self[name] = self._store[name]

    def close(self):
        pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


class FakeIndex(object):
    def __init__(self, chrom, copy, copy_index, chrom_sizes):
        self.chrom = np.asarray(chrom)
        self.copy = np.asarray(copy)
        self.copy_index = copy_index          # dict haploid bin -> list of bead ids
        self.chrom_sizes = np.asarray(chrom_sizes)

    def __len__(self):
        return len(self.chrom)


class FakeHss(object):
    """Duck-typed alabtools.analysis.HssFile over in-RAM arrays
    (methods used by get_actdist: ActivationDistanceStep.py:359-393)."""

    def __init__(self, coordinates, radii, index):
        self.coordinates = coordinates        # (nbead, nstruct, 3) float32
        self.radii = radii
        self.index = index

    def get_nstruct(self):
        return self.coordinates.shape[1]

    def get_index(self):
        return self.index

    def get_radii(self):
        return self.radii

    def get_bead_crd(self, b):
        return self.coordinates[b]

    def get_struct_crd(self, s):
        return self.coordinates[:, s, :]

    def close(self):
        pass


def import_reference():
    if "igm" in sys.modules:
        return sys.modules["igm"]
    h5py = types.ModuleType("h5py")
    h5py.File = _FakeH5File
    sys.modules["h5py"] = h5py
    alab = types.ModuleType("alabtools")
    alab.Contactmatrix = type("Contactmatrix", (), {})
    ana = types.ModuleType("alabtools.analysis")
    ana.HssFile = FakeHss
    ana.COORD_DTYPE = np.float32
    utils = types.ModuleType("alabtools.utils")
    for n in ("Genome", "Index", "make_diploid", "make_multiploid"):
        setattr(utils, n, type(n, (), {}))
    plots = types.ModuleType("alabtools.plots")
    plots.plot_comparison = lambda *a, **k: None
    alab.analysis, alab.utils, alab.plots = ana, utils, plots
    sys.modules.update({"alabtools": alab, "alabtools.analysis": ana,
                        "alabtools.utils": utils, "alabtools.plots": plots})
    mpl = types.ModuleType("matplotlib")
    plt = types.ModuleType("matplotlib.pyplot")
    mpl.pyplot = plt
    mpl.use = lambda *a, **k: None
    sys.modules.update({"matplotlib": mpl, "matplotlib.pyplot": plt})
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    # the reference's own native SPRITE extension, compiled by oracle/build_ref.sh
    import glob
    import importlib.util
    import os
    so = glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)),
                                "..", "..", "oracle", "_ref", "sprite*.so"))
    if not so:
        raise RuntimeError("run oracle/build_ref.sh first")
    pkg = types.ModuleType("igm.cython_compiled")
    pkg.__path__ = []
    spec = importlib.util.spec_from_file_location("igm.cython_compiled.sprite", so[0])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["igm.cython_compiled"] = pkg
    sys.modules["igm.cython_compiled.sprite"] = mod
    spec.loader.exec_module(mod)
    pkg.sprite = mod
    import igm
    igm.cython_compiled = pkg
    return igm


def import_staged(root):
    """Import the staged pure-Python part of the reference (igm.model, igm.restraints; oracle/build_ref.sh copies it to
    oracle/_ref/igm_py) with the h5py stand-in above -- what the GPU box has instead of /root/reference."""
    import os
    if "igm" in sys.modules:
        return sys.modules["igm"]
    h5py = types.ModuleType("h5py")
    h5py.File = _FakeH5File
    sys.modules["h5py"] = h5py
    if root not in sys.path:
        sys.path.insert(0, root)
    import igm.model          # noqa: F401
    import igm.restraints     # noqa: F401
    return sys.modules["igm"]
