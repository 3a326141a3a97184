"""Bindings with the shape of the reference's A-step tasks, so that the UNCHANGED ``ActivationDistanceStep`` (setup,
batching, reduce, file protocol) runs its map phase on the GPU:

    from igm.steps import ActivationDistanceStep
    from igm_b200 import steps_adapter
    ActivationDistanceStep.task = staticmethod(steps_adapter.actdist_task)

``actdist_task(batch_id, cfg, tmp_dir)`` does what ``ActivationDistanceStep.task`` does
(igm/steps/ActivationDistanceStep.py:192-222): reads ``<tmp_dir>/<batch_id>.in.npy`` (rows of i, j, pwish, plast, :179),
computes the rows of ``get_actdist`` (:319-459) for every pair of the batch over the population in
``cfg['optimization']['structure_output']`` and writes ``<tmp_dir>/<batch_id>.out.tmp`` in the reference's text format
(:39, :219-222), which ``reduce`` (:238-249) parses back into exactly the float32 values the kernel produced.  The
population file is read once per process and stays resident on the GPU (the reference re-opens it in every task).
"""
import os
import threading

import numpy as np

from . import files
from .astep import Population
from .kernel import _cfg_get

ACTDIST_FMT = "%6d %6d %10.4f %.4f"          # actdist_fmt_str, ActivationDistanceStep.py:39

_lock = threading.Lock()
_resident = {}                               # (path, mtime, device) -> Population


def resident_population(path, device=0):
    """the population of an .hss file on the GPU, loaded on first use and whenever the file changed"""
    key = (os.path.realpath(path), os.path.getmtime(path), int(device))
    with _lock:
        pop = _resident.get(key)
        if pop is None:
            for k in [k for k in _resident if k[0] == key[0] and k[2] == key[2]]:
                _resident.pop(k).close()                     # an older version of the same file
            hss = files.read_hss(path)
            idx = hss["index"]
            pop = Population(hss["nstruct"], hss["radii"], idx.chrom, idx.copy_off, idx.copy_beads, device=device)
            pop.upload(np.ascontiguousarray(hss["coordinates"], dtype=np.float32))
            _resident[key] = pop
        return pop


def release():
    with _lock:
        for pop in _resident.values():
            pop.close()
        _resident.clear()


def actdist_task(batch_id, cfg, tmp_dir):
    """Drop-in for ``ActivationDistanceStep.task`` (same arguments, same files in and out, returns None)."""
    contact_range = _cfg_get(cfg, "restraints/Hi-C/contact_range", 2.0)
    device = _cfg_get(cfg, "optimization/kernel_opts/b200/device", 0) or 0
    pop = resident_population(_cfg_get(cfg, "optimization/structure_output"), device)
    params = np.load(os.path.join(tmp_dir, "%d.in.npy" % batch_id))
    params = np.asarray(params, dtype=np.float64).reshape(-1, 4)                    # an empty batch is saved as shape (0,)
    with _lock:                                                                     # one pair batch at a time per population
        row, col, dist, prob = pop.actdist(params[:, 0].astype(np.int32), params[:, 1].astype(np.int32), params[:, 2], params[:, 3],
                                           contact_range)
    with open(os.path.join(tmp_dir, "%d.out.tmp" % batch_id), "w") as f:
        f.write("\n".join([ACTDIST_FMT % x for x in zip(row.tolist(), col.tolist(), dist.tolist(), prob.tolist())]))
