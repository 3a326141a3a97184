"""Population files without h5py / alabtools (SURVEY §8 f rank 2).  The reader is checked against the
reference's own h5py-written demo files when the reference tree is present (this container), the writer by
reading back what it wrote."""
import json
import os

import numpy as np
import pytest

from igm_b200 import files, h5lite

DEMO = "/root/reference/demo"
needs_demo = pytest.mark.skipif(not os.path.isdir(DEMO), reason="reference demo files not present")


def _same(a, b, path=""):
    assert set(a) == set(b), (path, set(a) ^ set(b))
    for k in a:
        if isinstance(a[k], dict):
            _same(a[k], b[k], path + "/" + k)
        elif isinstance(a[k], str):
            assert a[k] == b[k], (path, k)
        else:
            x, y = np.asarray(a[k]), np.asarray(b[k])
            if x.dtype.kind in "USO":
                assert list(x.ravel()) == list(y.ravel()), (path, k)
            else:
                assert x.dtype == y.dtype and x.shape == y.shape and np.array_equal(x, y), (path, k)


def test_write_then_read_every_supported_kind(tmp_path):
    rng = np.random.default_rng(0)
    tree = {"coordinates": rng.normal(size=(7, 5, 3)).astype(np.float32), "radii": np.full(7, 280.7, np.float32),
            "summary": '{"a": 1, "b": "é"}', "empty": np.zeros(0, np.float64), "f8": np.float64(3.25),
            "u1": np.arange(5, dtype=np.uint8), "i8": np.arange(4, dtype=np.int64).reshape(2, 2),
            "index": {"chromstr": np.array(["chr1", "chr1", "chr10"]), "copy": np.arange(3, dtype=np.int32),
                      "copy_index": '{"0": [0, 3]}', "deeper": {"x": np.float32(1.5)}},
            h5lite.ATTRS: {"nbead": np.int64(7), "nstruct": np.int64(5), "violation": np.float64(1e-6), "version": np.int32(2)}}
    p = str(tmp_path / "t.h5")
    h5lite.write(p, tree)
    _same(tree, h5lite.read_tree(p))
    with open(p, "rb") as fh:
        head = fh.read(96)
    assert head[:8] == b"\x89HDF\r\n\x1a\n" and head[8] == 0 and head[13] == 8 and head[14] == 8     # superblock v0, 8-byte offsets
    assert int.from_bytes(head[40:48], "little") == os.path.getsize(p)                                 # end-of-file address


def test_actdist_file_roundtrip(tmp_path):
    rng = np.random.default_rng(1)
    row = rng.integers(0, 3008, 1000).astype(np.int32); col = rng.integers(0, 3008, 1000).astype(np.int32)
    dist = rng.uniform(100, 2000, 1000).astype(np.float32); prob = rng.uniform(0, 1, 1000).astype(np.float32)
    p = str(tmp_path / "actdist.hdf5")
    files.write_actdist(p, row, col, dist, prob)
    got = files.read_actdist(p)
    assert all(np.array_equal(a, b) and a.dtype == b.dtype for a, b in zip(got, (row, col, dist, prob)))   # :277-281 dtypes
    files.write_actdist(p, row[:0], col[:0], dist[:0], prob[:0])                                          # no pair kept
    assert all(a.size == 0 for a in files.read_actdist(p))


@needs_demo
def test_reads_the_reference_population_and_matrix():
    pop = files.read_hss(os.path.join(DEMO, "demo_sample_outputs", "igm-model.hss.T"))
    assert pop["coordinates"].shape == (3008, 100, 3) and pop["nstruct"] == 100 and pop["nbead"] == 3008
    assert np.allclose(pop["radii"], 280.7308) and pop["index"].n_hap == 1558
    assert pop["index"].copy_index[0] == [0, 1558] and pop["index"].copy_off[-1] == 3008
    assert json.loads(pop["tree"]["summary"])["n_imposed"] > 1e6
    hcs = files.read_hcs(os.path.join(DEMO, "WTC11_HiC_2Mb.hcs"))
    assert hcs["n"] == 1558 and hcs["indptr"][-1] == hcs["data"].size == hcs["indices"].size
    # pairs kept per sigma: the counts the reference's own run log shows for this matrix (SURVEY.md §8 d)
    assert files.hic_pairs_from_matrix(hcs, 0.01, 0.01)[0].size == 614140
    pi, pj, pw, pl = files.hic_pairs_from_matrix(hcs, 0.2, False)
    ch = hcs["index"].chrom
    assert pi.size > 1000 and (ch[pi] == ch[pj]).all() and (pw >= np.float32(0.2)).all() and (pi < pj).all() and not pl.any()
    assert np.array_equal(np.lexsort((pj, pi)), np.arange(pi.size))                    # CSR (row-major) order


@needs_demo
def test_rewrites_the_reference_population_identically(tmp_path):
    src = os.path.join(DEMO, "demo_sample_outputs", "igm-model.hss.T")
    pop = files.read_hss(src)
    p = str(tmp_path / "copy.hss")
    files.write_hss(p, pop["tree"])
    _same(pop["tree"], h5lite.read_tree(p))
    new = pop["coordinates"][:, :7, :] + np.float32(1.0)
    files.write_hss(p, pop["tree"], coordinates=new, summary={"n_imposed": 3}, violation=0.25)
    back = files.read_hss(p)
    assert back["nstruct"] == 7 and np.array_equal(back["coordinates"], new)
    assert back["tree"][h5lite.ATTRS]["violation"] == 0.25 and json.loads(back["tree"]["summary"]) == {"n_imposed": 3}


def test_plast_lookup_sums_duplicates_and_ignores_second_copies():
    hcs = {"n": 4, "indptr": np.array([0, 2, 3, 3, 3], dtype=np.int32), "indices": np.array([1, 3, 2], dtype=np.int32),
           "data": np.array([0.5, 0.05, 0.3], dtype=np.float32), "diagonal": np.zeros(4, np.float32),
           "index": files.Index([0, 0, 1, 1, 0, 0], [0, 0, 0, 0, 1, 1], {0: [0, 4], 1: [1, 5], 2: [2], 3: [3]})}
    last = (np.array([0, 4, 0, 1], dtype=np.int32), np.array([1, 5, 1, 2], dtype=np.int32), np.zeros(4, np.float32),
            np.array([0.125, 0.5, 0.25, 0.0625], dtype=np.float32))
    pi, pj, pw, pl = files.hic_pairs_from_matrix(hcs, 0.1, 0.1, last)
    assert pi.tolist() == [0, 1] and pj.tolist() == [1, 2] and pw.tolist() == [0.5, float(np.float32(0.3))]
    assert pl.tolist() == [0.375, 0.0625]            # (0,1) twice among first copies: summed (coo -> lil); (4,5) ignored
    assert files.hic_pairs_from_matrix(hcs, False, 0.01)[0].tolist() == [0, 1]          # inter only: (0,3) and (1,2)


@needs_demo
def test_recorded_demo_output_counts_are_the_uncorrected_probability():
    """the Hi-C restraint counts in the reference's recorded demo output (sigma 0.02) equal, to 1.5 %,
    the sum over the kept pairs of round(pwish n S) + 1 -- what p = pwish imposes -- per class."""
    pop = files.read_hss(os.path.join(DEMO, "demo_sample_outputs", "igm-model.hss.T"))
    cfg = json.loads(pop["tree"]["config_data"])
    assert cfg["runtime"]["Hi-C"]["intra_sigma"] == 0.02 and cfg["runtime"]["Hi-C"]["inter_sigma"] == 0.02
    rec = {k.split("[")[0]: v["n_imposed"] for k, v in json.loads(pop["tree"]["summary"])["byrestraint"].items()}
    hcs = files.read_hcs(os.path.join(DEMO, "WTC11_HiC_2Mb.hcs"))
    pi, pj, pw, _ = files.hic_pairs_from_matrix(hcs, 0.02, 0.02)
    nc = np.diff(pop["index"].copy_off)
    same = hcs["index"].chrom[pi] == hcs["index"].chrom[pj]
    ncomb = np.where(same, np.minimum(nc[pi], nc[pj]), nc[pi] * nc[pj])
    S = pop["nstruct"]
    o = np.minimum(ncomb * S - 1, np.rint(ncomb * pw * S)) + 1
    assert abs(o[same].sum() / rec["intraHiC"] - 1) < 0.015 and abs(o[~same].sum() / rec["interHiC"] - 1) < 0.02
    assert rec["Envelope"] == pop["nbead"] * S and rec["Polymer"] == 2962 * S
