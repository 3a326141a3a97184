"""Synthetic inputs of the shapes BASELINE.json names (there is no network for real data
on the GPU box).  Everything is seeded and reproducible.

Geometry follows the reference's demo: hg38 male at 2 Mb = 1558 haploid
bins (1450 autosomal, 79 X, 29 Y) -> 3008 beads in 46 chains, bead order = first copies of
chr1..22,X,Y followed by the second copies of chr1..22 (what alabtools' make_diploid yields for
igm/_preprocess.py:47-74); radii from the occupancy formula of igm/_preprocess.py:128-131;
start coordinates from ``generate_territories`` (igm/steps/RandomInit.py:69-102).
"""
import math

import numpy as np

from ._abi import PackedSystem, KIND_BEAD, KIND_STATIC, BOND_UPPER

# 2 Mb bins per chromosome of the demo index (chr1..22, X, Y)
CHROM_BINS_2MB = [125, 122, 100, 96, 91, 86, 80, 73, 70, 67, 68, 67, 58, 54, 51, 46, 42, 41, 30, 33, 24, 26, 79, 29]
# hg38 chromosome lengths / 200 kb, rounded up (29 838 beads diploid male, IGM_documentation.pdf)
CHROM_BP_HG38 = [248956422, 242193529, 198295559, 190214555, 181538259, 170805979, 159345973, 145138636,
                 138394717, 133797422, 135086622, 133275309, 114364328, 107043718, 101991189, 90338345,
                 83257441, 80373285, 58617616, 64444167, 46709983, 50818468, 156040895, 57227415]


class Genome(object):
    """What the hot path reads from alabtools' Index: chrom, copy, copy_index, chrom_sizes."""

    def __init__(self, chrom_bins, n_autosomes=22):
        chrom, copy = [], []
        for c, nb in enumerate(chrom_bins):
            chrom += [c] * nb
            copy += [0] * nb
        n_hap = len(chrom)
        second = []
        for c in range(n_autosomes):
            second += [c] * chrom_bins[c]
        chrom += second
        copy += [1] * len(second)
        self.chrom = np.array(chrom, dtype=np.int32)
        self.copy = np.array(copy, dtype=np.int32)
        self.n_hap = n_hap
        self.n_bead = len(chrom)
        self.chrom_sizes = np.array(list(chrom_bins) + list(chrom_bins[:n_autosomes]), dtype=np.int32)
        # copy_index: haploid bin -> bead ids of its copies
        start_hap = np.concatenate([[0], np.cumsum(chrom_bins)])
        off2 = n_hap
        self.copy_index = {}
        for c, nb in enumerate(chrom_bins):
            for b in range(nb):
                h = int(start_hap[c] + b)
                self.copy_index[h] = [h] if c >= n_autosomes else [h, int(off2 + start_hap[c] + b)]
        off = np.zeros(n_hap + 1, dtype=np.int32)
        beads = []
        for h in range(n_hap):
            beads += self.copy_index[h]
            off[h + 1] = len(beads)
        self.copy_off = off
        self.copy_beads = np.array(beads, dtype=np.int32)

    def __len__(self):
        return self.n_bead


def genome_2mb():
    return Genome(CHROM_BINS_2MB)


def genome_200kb():
    return Genome([int(math.ceil(bp / 200000.0)) for bp in CHROM_BP_HG38])


def bead_radii(n_bead, nucleus_radius=5500.0, occupancy=0.4):
    """igm/_preprocess.py:128-131 with equal-sized bins: occupancy * V_nucleus shared equally."""
    r = nucleus_radius * (occupancy / float(n_bead)) ** (1.0 / 3.0)
    return np.full(n_bead, r, dtype=np.float32)


def _uniform_sphere(rng, R, n):
    phi = rng.uniform(0, 2 * math.pi, n)
    costheta = rng.uniform(-1, 1, n)
    u = rng.uniform(0, 1, n)
    theta = np.arccos(costheta)
    r = R * u ** (1.0 / 3.0)
    return np.stack([r * np.sin(theta) * np.cos(phi), r * np.sin(theta) * np.sin(phi), r * np.cos(theta)], axis=1)


def territories(genome, R, rng):
    """generate_territories (igm/steps/RandomInit.py:69-102), vectorised per chain."""
    n_tot = len(genome)
    chr_radii = [0.75 * R * (float(nb) / n_tot) ** (1.0 / 3) for nb in genome.chrom_sizes]
    crad = float(np.average(chr_radii))
    out = np.empty((n_tot, 3))
    k = 0
    for nb in genome.chrom_sizes:
        center = _uniform_sphere(rng, R - crad, 1)[0]
        out[k:k + nb] = _uniform_sphere(rng, crad, int(nb)) + center
        k += nb
    return out.astype(np.float32)


def population_territories(genome, n_struct, R=5500.0, seed=1234):
    """[n_bead][n_struct][3] float32 (the .hss layout)."""
    crd = np.empty((len(genome), n_struct, 3), dtype=np.float32)
    for s in range(n_struct):
        crd[:, s, :] = territories(genome, R, np.random.default_rng(seed + s))
    return crd


def polymer_bonds(genome, radii, contact_range=2.0):
    """Polymer._apply (igm/restraints/polymer.py:31-55), 'simple' style."""
    i = np.arange(len(genome) - 1)
    ok = (genome.chrom[i] == genome.chrom[i + 1]) & (genome.copy[i] == genome.copy[i + 1])
    i = i[ok].astype(np.int32)
    r0 = contact_range * (radii[i].astype(np.float32) + radii[i + 1].astype(np.float32))
    return i, i + 1, r0.astype(np.float64)


def hic_pairs(genome, sigma, seed=7, intra_scale=0.9, intra_exp=1.08, inter_frac=0.496, inter_scale=0.0024):
    """Synthetic haploid Hi-C probabilities thresholded at ``sigma`` in the order
    Contactmatrix.coo_generator yields them (row-major upper triangle,
    ActivationDistanceStep.py:167-186): p = min(1, 0.9 |i-j|^-1.08) within a chromosome; between
    chromosomes a fraction ``inter_frac`` of the pairs gets 0.01 + Exp(inter_scale) and the rest
    U(0, 0.01), which reproduces the demo matrix' pair counts per sigma (SURVEY.md section 8:
    ~54 k intra / ~560 k inter at 0.01, ~8 k inter at 0.02).
    -> (pair_i i4, pair_j i4, pwish f8 holding float32 values)."""
    rng = np.random.default_rng(seed)
    nh = genome.n_hap
    ch = genome.chrom[:nh]
    pi, pj, pw = [], [], []
    for i in range(nh - 1):
        j = np.arange(i + 1, nh)
        same = ch[j] == ch[i]
        p = np.empty(j.size, dtype=np.float32)
        d = (j[same] - i).astype(np.float64)
        p[same] = np.minimum(1.0, intra_scale * d ** (-intra_exp)).astype(np.float32)
        m = int((~same).sum())
        hi = rng.random(m) < inter_frac
        p[~same] = np.where(hi, 0.01 + rng.exponential(inter_scale, m), rng.uniform(0.0, 0.01, m)).astype(np.float32)
        keep = p >= np.float32(sigma)
        pi.append(np.full(int(keep.sum()), i, dtype=np.int32))
        pj.append(j[keep].astype(np.int32))
        pw.append(p[keep].astype(np.float64))
    return np.concatenate(pi), np.concatenate(pj), np.concatenate(pw)


def pack_structure(xyz, radii, genome, hic_i=None, hic_j=None, hic_r0=None, nucleus_radius=5500.0,
                   semiaxes=None, contact_range=2.0, evfactor=1.0, k_poly=1.0, k_hic=1.0, k_env=1.0,
                   seed=1, struct_id=0, extra_envelopes=(), n_extra_dummies=0):
    """One structure the way ModelingStep.task assembles it (igm/steps/ModelingStep.py:184-264):
    beads, Steric, Envelope (adds its DUMMY_STATIC centre particle, envelope.py:39), Polymer and
    the assigned Hi-C HarmonicUpperBounds.  n_extra_dummies: further DUMMY_STATIC particles at the origin,
    the centres the Damid and Fish restraints add (damid.py:120, fish.py:119), atoms n+1, n+2, ..."""
    n = len(genome)
    nd = 1 + int(n_extra_dummies)
    xyz_all = np.vstack([xyz.astype(np.float32), np.zeros((nd, 3), dtype=np.float32)])
    rad_all = np.concatenate([radii.astype(np.float32), np.zeros(nd, dtype=np.float32)])
    kind = np.concatenate([np.full(n, KIND_BEAD, dtype=np.uint8), np.full(nd, KIND_STATIC, dtype=np.uint8)])
    bi, bj, r0 = polymer_bonds(genome, radii, contact_range)
    bk = np.full(bi.size, k_poly, dtype=np.float64)
    if hic_i is not None and len(hic_i):
        bi = np.concatenate([bi, np.asarray(hic_i, dtype=np.int32)])
        bj = np.concatenate([bj, np.asarray(hic_j, dtype=np.int32)])
        r0 = np.concatenate([r0, np.asarray(hic_r0, dtype=np.float64)])
        bk = np.concatenate([bk, np.full(len(hic_i), k_hic, dtype=np.float64)])
    if semiaxes is None:
        semiaxes = (nucleus_radius,) * 3
    env_abck = [tuple(semiaxes) + (k_env,)]
    env_members = [np.arange(n, dtype=np.int32)]
    for (abc, k, members) in extra_envelopes:
        env_abck.append(tuple(abc) + (k,))
        env_members.append(np.asarray(members, dtype=np.int32))
    env_off = np.concatenate([[0], np.cumsum([m.size for m in env_members])]).astype(np.int32)
    return PackedSystem(xyz_all, rad_all, kind, bi, bj, bk, r0, np.full(bi.size, BOND_UPPER, dtype=np.uint8),
                        np.array(env_abck, dtype=np.float64), env_off, np.concatenate(env_members),
                        evfactor, seed=seed, struct_id=struct_id)


def demo_protocol_options():
    """optimization/optimizer_options of demo/config_file.json:80-131 (the protocol every
    BASELINE config uses)."""
    return {
        "mdsteps": 45000, "timestep": 0.25, "tstart": 500.0, "tstop": 0.01,
        "custom_annealing_protocol": {
            "num_steps": 4,
            "mdsteps": [5000, 15000, 15000, 10000],
            "tstarts": [5000.0, 500.0, 50.0, 1.0],
            "tstops": [500.0, 50.0, 1.0, 0.0],
            "evfactors": [0.5, 1.0, 1.0, 1.0],
            "envelope_factors": [1.2, 1.0, 1.0, 1.0],
            "relax": {"mdsteps": 500, "temperature": 1.0, "max_velocity": 10.0},
        },
        "damp": 50.0, "max_velocity": 1000.0, "etol": 0.0001, "ftol": 1e-06,
        "max_cg_iter": 500, "max_cg_eval": 500, "thermo": 1000, "write": -1,
    }
