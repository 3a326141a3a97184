"""Host helpers for ``igmk_single_point_noalloc``, the single-point entry for callers that may not
allocate GPU memory (a service running on one pre-allocated arena, a latency-critical loop that cannot
afford the implicit synchronisation of cudaMalloc/cudaFree, a card another tenant has filled).  The
library keeps the kernel's scratch in a fixed array inside its own image, so the call itself
allocates nothing; the CALLER owns every input and output buffer on the device.  This module only
builds the flat host arrays in the layout the entry point expects and wraps the ctypes call; how the
arrays get into the caller's device memory is the caller's business.
"""
import ctypes as C

import numpy as np

from . import _abi
from ._abi import KIND_BEAD, BOND_LOWER

MAX_ATOMS = 3328            # IGMK_NOALLOC_MAX_ATOMS in include/igmk.h


def host_layout(ps):
    """PackedSystem -> dict of host arrays in the device layouts documented in include/igmk.h."""
    n = ps.n_atoms
    xyzr = np.empty((n, 4), dtype=np.float32)
    xyzr[:, :3] = ps.xyz
    xyzr[:, 3] = ps.radius
    envmask = np.zeros(n, dtype=np.uint32)
    for e in range(ps.env_abck.shape[0]):
        envmask[ps.env_members[ps.env_off[e]:ps.env_off[e + 1]]] |= np.uint32(1 << e)
    # bond types (r0, +-k), de-duplicated; every bond listed under both of its atoms
    ks = np.where(ps.bond_style == BOND_LOWER, -ps.bond_k, ps.bond_k)
    types, tidx = np.unique(np.stack([ps.bond_r0, ks], axis=1), axis=0, return_inverse=True) if ps.bond_i.size \
        else (np.zeros((1, 2)), np.zeros(0, dtype=np.int64))
    tidx = np.asarray(tidx).reshape(-1)
    if types.shape[0] > 65536:
        raise ValueError("more than 65536 distinct bond types: the item format (partner | type << 16) cannot hold them")
    if n > 65535:
        raise ValueError("more than 65535 atoms: the item format (partner | type << 16) cannot hold them")
    owner = np.concatenate([ps.bond_i, ps.bond_j]).astype(np.int64)
    partner = np.concatenate([ps.bond_j, ps.bond_i]).astype(np.uint32)
    tt = np.concatenate([tidx, tidx]).astype(np.uint32)
    order = np.argsort(owner, kind="stable")
    bond_ptr = np.zeros(n + 1, dtype=np.int32)
    np.cumsum(np.bincount(owner, minlength=n), out=bond_ptr[1:])
    items = (partner[order] | (tt[order] << np.uint32(16))).astype(np.uint32)
    if items.size == 0:
        items = np.zeros(1, dtype=np.uint32)
    beads = ps.kind == KIND_BEAD
    return {
        "xyzr": xyzr, "kind": ps.kind.copy(), "envmask": envmask, "bond_ptr": bond_ptr, "bond_items": items,
        "btype": np.ascontiguousarray(types, dtype=np.float64), "env_abck": np.ascontiguousarray(ps.env_abck, dtype=np.float64),
        "evfactor": ps.evfactor, "max_radius": float(ps.radius[beads].max()) if beads.any() else 0.0,
    }


def single_point_noalloc(device, stream, n_atoms, d_xyzr, d_kind, d_envmask, d_bond_ptr, d_bond_items, d_btype,
                         env_abck, evfactor, max_radius, d_forces, d_energies, d_status,
                         ev_scale=1.0, env_scale=1.0, bond_prefactor=1.0):
    """Thin wrapper: the ``d_*`` arguments are integer device addresses of caller-owned buffers.
    Asynchronous on ``stream`` (integer cudaStream_t handle or 0); synchronise before reading the outputs."""
    env = np.ascontiguousarray(env_abck, dtype=np.float64).reshape(-1, 4)
    _abi.check(_abi.load().igmk_single_point_noalloc(
        int(device), C.c_void_p(stream or None), int(n_atoms), C.c_void_p(d_xyzr), C.c_void_p(d_kind), C.c_void_p(d_envmask),
        C.c_void_p(d_bond_ptr), C.c_void_p(d_bond_items), C.c_void_p(d_btype), env.shape[0],
        env.ctypes.data_as(_abi.c_f64p), float(evfactor), float(max_radius), float(ev_scale), float(env_scale),
        float(bond_prefactor), C.c_void_p(d_forces), C.c_void_p(d_energies), C.c_void_p(d_status)),
        "igmk_single_point_noalloc")
