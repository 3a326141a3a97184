"""CPU restatement (numpy) of IGM's Hi-C A-step.  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs; the product path never touches it.

Pinned: tests/golden/astep_golden.npz holds outputs of the reference's own
``get_actdist`` and ``inter/intraHiC._apply`` run in the build container on the
demo population (tests/golden/make_golden.py); tests/test_oracle_astep.py checks
this restatement against them bit for bit.

Numeric environment restated: numpy >= 2 scalar promotion (float32 scalar * Python
float stays float32) and the x86-64 OpenBLAS ``sdot`` tail loop
behind ``np.linalg.norm`` of a 3-vector (float32 products, double accumulator,
SURVEY.md 8(a) A5).
"""
import numpy as np

ACTDIST_FMT = "%6d %6d %10.4f %.4f"      # ActivationDistanceStep.py:39


def clean_probability(pij, pexist):
    """ActivationDistanceStep.py:306-315"""
    if pexist < 1:
        pclean = (pij - pexist) / (1.0 - pexist)
    else:
        pclean = pij
    return max(0, pclean)


def text_roundtrip(x, ndec=4):
    """'%10.4f' / '%.4f' formatting (ActivationDistanceStep.py:39,221) followed by
    np.genfromtxt's float32 field parse (:241): float64(str) then cast to float32."""
    return np.float32(float(("%." + str(ndec) + "f") % x))


def get_actdist(i, j, pwish, plast, crd, radii, chrom, copy_index, contact_range=2.0):
    """ActivationDistanceStep.py:319-459 for one haploid pair.

    crd: (nbead, nstruct, 3) float32 (hss 'coordinates'); copy_index: haploid bin ->
    list of bead ids.  Returns [(row, col, dist_f64, p_f64), ...] like the reference
    (before the text round trip)."""
    if i == j:
        return []                                               # :351-352
    n_struct = crd.shape[1]
    ii, jj = copy_index[i], copy_index[j]                       # :361-362
    ri, rj = radii[ii[0]], radii[jj[0]]                         # :365 (np.float32 scalars)
    rcutsq = np.square(np.float32(contact_range) * (ri + rj))   # :368, float32 under numpy 2
    if chrom[i] == chrom[j]:                                    # :377-393 intra: zip(ii, jj)
        pairs = list(zip(ii, jj))
        n_poss = min(len(ii), len(jj))
        n_comb = len(ii)
    else:                                                       # :395-413 inter: ii x jj
        pairs = [(k, m) for k in ii for m in jj]
        n_poss = n_comb = len(ii) * len(jj)
    d_sq = np.empty((n_comb, n_struct))                         # float64 container
    for it, (k, m) in enumerate(pairs):
        d = crd[k] - crd[m]                                     # float32
        sq = d * d
        d_sq[it] = (sq[:, 0] + sq[:, 1]) + sq[:, 2]             # np.sum(axis=1) == (a+b)+c in float32
    d_sq = d_sq[:n_poss]
    contact_count = np.count_nonzero(d_sq <= rcutsq)            # :419
    pnow = float(contact_count) / (n_poss * n_struct)           # :422
    sortdist_sq = np.sort(d_sq.ravel())                         # :425
    t = clean_probability(pnow, plast)                          # :430
    p = clean_probability(pwish, t)                             # :434
    res = []
    if p > 0:
        o = min(n_poss * n_struct - 1, int(round(n_poss * p * n_struct)))   # :443-444
        ad = np.sqrt(sortdist_sq[o])                            # :446 (float64 sqrt)
        if chrom[i] == chrom[j]:
            res = [(i0, i1, ad, p) for i0, i1 in zip(ii, jj)]   # :450-451
        else:
            res = [(i0, i1, ad, p) for i0 in ii for i1 in jj]   # :456
    return res


def actdist_rows(pairs_i, pairs_j, pwish, plast, crd, radii, chrom, copy_index, contact_range=2.0):
    """task() + reduce() of ActivationDistanceStep (:192-222, :238-249): all rows of a
    batch after the text round trip -> (row i4, col i4, dist f4, prob f4)."""
    row, col, dist, prob = [], [], [], []
    for i, j, pw, pl in zip(pairs_i, pairs_j, pwish, plast):
        for (a, b, ad, p) in get_actdist(int(i), int(j), float(pw), float(pl), crd, radii, chrom,
                                         copy_index, contact_range):
            row.append(a)
            col.append(b)
            dist.append(text_roundtrip(ad, 4))
            prob.append(text_roundtrip(p, 4))
    return (np.array(row, dtype=np.int32), np.array(col, dtype=np.int32),
            np.array(dist, dtype=np.float32), np.array(prob, dtype=np.float32))


def particle_distance(xi, xj):
    """Particle.__sub__ (particle.py:35-36) = np.linalg.norm of a float32 3-vector
    = sqrt(sdot(d, d)): float32 products, double accumulation in index order, one
    rounding to float32, float32 sqrt.  Vectorised over leading axes."""
    d = (xi - xj).astype(np.float32)
    sq = (d * d).astype(np.float32)
    acc = (sq[..., 0].astype(np.float64) + sq[..., 1].astype(np.float64)) + sq[..., 2].astype(np.float64)
    return np.sqrt(acc.astype(np.float32))


def assign(crd, row, col, dist):
    """Contact-assignment predicate of interHiC/intraHiC._apply (inter_hic.py:53,
    intra_hic.py:54) for every (actdist row, structure): bool [n_rows, n_struct].
    (The intra/inter chrom test only routes a bond to one of the two restraint
    objects; every row lands in exactly one of them.)"""
    out = np.zeros((len(row), crd.shape[1]), dtype=bool)
    for r in range(len(row)):
        out[r] = particle_distance(crd[row[r]], crd[col[r]]) <= dist[r]
    return out


def pack_bits(mask):
    """bool [n_rows, n_struct] -> uint32 [n_rows, ceil(n_struct/32)] (bit s%32 of word s//32)."""
    n_rows, n_struct = mask.shape
    nw = (n_struct + 31) // 32
    padded = np.zeros((n_rows, nw * 32), dtype=np.uint8)
    padded[:, :n_struct] = mask
    b = padded.reshape(n_rows, nw, 32).astype(np.uint32)
    return (b << np.arange(32, dtype=np.uint32)).sum(axis=2, dtype=np.uint64).astype(np.uint32)
