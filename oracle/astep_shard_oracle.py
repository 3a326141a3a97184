"""numpy stand-in for one structure shard of the sharded Hi-C A-step (igm_b200/parallel.py: sharded_actdist).
TEST INFRASTRUCTURE ONLY: lets the multi-rank protocol (counts and radix histograms all-reduced over ranks) run on
CPU under gloo; the arithmetic is get_actdist's (ActivationDistanceStep.py:319-459) as restated in astep_oracle.py."""
import numpy as np

from .astep_oracle import clean_probability, text_roundtrip


class NumpyShard(object):
    def __init__(self, crd, radii, chrom, copy_index):
        self.crd, self.radii, self.chrom, self.ci = crd, radii, chrom, copy_index      # crd: this shard's [n_bead][S_local][3]

    def _pairs_of(self, i, j):
        ii, jj = self.ci[int(i)], self.ci[int(j)]
        if self.chrom[i] == self.chrom[j]:
            return list(zip(ii, jj))
        return [(k, m) for k in ii for m in jj]

    def _dsq(self, i, j):
        out = []
        for k, m in self._pairs_of(i, j):
            d = self.crd[k] - self.crd[m]
            sq = d * d
            out.append(((sq[:, 0] + sq[:, 1]) + sq[:, 2]).astype(np.float32))
        return np.concatenate(out) if out else np.zeros(0, dtype=np.float32)

    def set_pairs(self, pi, pj, pw, pl):
        self.pi, self.pj, self.pw, self.pl = pi, pj, pw, pl
        self.m = pi.size
        self.keys = [self._dsq(i, j).view(np.uint32) if i != j else np.zeros(0, np.uint32) for i, j in zip(pi, pj)]

    def count(self, contact_range):
        import torch
        c = np.zeros(self.m, dtype=np.int32)
        for q, (i, j) in enumerate(zip(self.pi, self.pj)):
            if i == j:
                continue
            ri, rj = self.radii[self.ci[int(i)][0]], self.radii[self.ci[int(j)][0]]
            rcutsq = np.square(np.float32(contact_range) * (ri + rj))
            c[q] = np.count_nonzero(self.keys[q].view(np.float32) <= rcutsq)
        return torch.from_numpy(c)

    def rank(self, n_struct_total, total_counts):
        tc = total_counts.numpy()
        self.valid = np.zeros(self.m, dtype=bool)
        self.want = np.zeros(self.m, dtype=np.int64)
        self.prefix = np.zeros(self.m, dtype=np.uint32)
        self.p = np.zeros(self.m)
        self.n_total = n_struct_total
        for q, (i, j) in enumerate(zip(self.pi, self.pj)):
            if i == j:
                continue
            ncomb = len(self._pairs_of(i, j))
            pnow = float(tc[q]) / (ncomb * n_struct_total)
            p = clean_probability(self.pw[q], clean_probability(pnow, self.pl[q]))
            if p > 0:
                self.valid[q] = True
                self.p[q] = p
                self.want[q] = min(ncomb * n_struct_total - 1, int(round(ncomb * p * n_struct_total)))

    def histogram(self, radix_pass):
        import torch
        shift = 24 - 8 * radix_pass
        h = np.zeros((self.m, 256), dtype=np.int32)
        for q in range(self.m):
            if not self.valid[q]:
                continue
            k = self.keys[q]
            if radix_pass > 0:
                hi = np.uint32((0xffffffff << (shift + 8)) & 0xffffffff)
                k = k[(k & hi) == self.prefix[q]]
            h[q] = np.bincount((k >> np.uint32(shift)) & np.uint32(0xff), minlength=256)
        return torch.from_numpy(h)

    def pick(self, radix_pass, total_hist):
        shift = 24 - 8 * radix_pass
        th = total_hist.numpy()
        for q in range(self.m):
            if not self.valid[q]:
                continue
            cum = np.cumsum(th[q].astype(np.int64))
            digit = int(np.searchsorted(cum, self.want[q], side="right"))
            self.want[q] -= (cum[digit - 1] if digit else 0)
            self.prefix[q] |= np.uint32(digit << shift)

    def sync(self):
        pass

    def collect(self):
        row, col, dist, prob = [], [], [], []
        for q, (i, j) in enumerate(zip(self.pi, self.pj)):
            if not self.valid[q]:
                continue
            ad = np.sqrt(np.float64(np.array([self.prefix[q]], dtype=np.uint32).view(np.float32)[0]))
            for a, b in self._pairs_of(i, j):
                row.append(a); col.append(b)
                dist.append(text_roundtrip(ad, 4)); prob.append(text_roundtrip(self.p[q], 4))
        return (np.array(row, dtype=np.int32), np.array(col, dtype=np.int32), np.array(dist, dtype=np.float32),
                np.array(prob, dtype=np.float32))
