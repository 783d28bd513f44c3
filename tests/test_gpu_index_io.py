"""GPU tests of the index file reader / writer: the reference's own "IwFl"
format (faiss/impl/index_write.cpp:635-638, index_read.cpp:782-787). Golden files under
tests/golden/*.ivfflat were written by the unmodified reference (tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest

from helpers import assert_knn_parity, clustered, compare_knn, sample_centroids

pytestmark = pytest.mark.gpu

GOLDEN_DIR = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("name", ["lowlevel_l2", "lowlevel_ip"])
def test_read_reference_file_and_search(cuda_device, name):
    """a file written by the reference loads onto the GPU: same fields, same lists in the same
    order, and the search returns what the reference returned"""
    import hnsw_ivf_b200 as b2
    from oracle import pyoracle as po
    g = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    path = os.path.join(GOLDEN_DIR, name + ".ivfflat")
    f = po.parse_index_file(path)
    idx = b2.read_index(path)
    assert (idx.d, idx.nlist, idx.nprobe, idx.ntotal, idx.metric_type, idx.is_trained) == (
        f["d"], f["nlist"], f["nprobe"], f["ntotal"], f["metric"], True)
    assert np.array_equal(idx.get_centroids(), g["centroids"])
    for l in range(idx.nlist):
        codes, ids = idx.get_list(l)
        assert np.array_equal(ids, f["lists"][l][1]) and np.array_equal(codes, f["lists"][l][0])
    k = int(g["k"])
    Dp, Ip = idx.search_preassigned(g["xq"], k, g["Iq"])
    assert_knn_parity(g["Dp"], g["Ip"], Dp, Ip, what="preassigned after read_index")
    D, I = idx.search(g["xq"], k)  # nprobe came from the file
    r = compare_knn(g["D"], g["I"], D, I)
    assert r["exact_id_rate"] > 0.99 and r["bad_dist"] == 0, r


@pytest.mark.parametrize("name", ["lowlevel_l2", "lowlevel_ip", "sparse_l2"])
def test_write_is_byte_identical_to_reference(cuda_device, tmp_path, name):
    """an index built on the device from the same inputs writes the same BYTES the reference wrote
    ("full" and "sprs" size tables)"""
    import hnsw_ivf_b200 as b2
    if name == "sparse_l2":
        g = np.load(os.path.join(GOLDEN_DIR, "sparse_l2.inputs"))
        metric, nlist, nprobe = 1, 40, 5
    else:
        g = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
        metric, nlist, nprobe = int(g["metric"]), int(g["nlist"]), int(g["nprobe"])
    idx = b2.IndexIVFFlat(g["xb"].shape[1], nlist, metric)
    idx.set_centroids(g["centroids"])
    idx.add_core(g["xb"], g["ids"], g["assign"])
    idx.nprobe = nprobe
    out = str(tmp_path / "mine.ivfflat")
    b2.write_index(idx, out)
    want = open(os.path.join(GOLDEN_DIR, name + ".ivfflat"), "rb").read()
    got = open(out, "rb").read()
    assert len(got) == len(want)
    assert got == want, "first differing byte at %d" % next(
        i for i in range(len(want)) if got[i] != want[i])
    # and the file reads back to the same index
    idx2 = b2.read_index(out)
    assert idx2.ntotal == idx.ntotal and idx2.nprobe == nprobe
    for l in range(nlist):
        a, b = idx.get_list(l), idx2.get_list(l)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


@pytest.mark.parametrize("metric", [1, 0])
def test_roundtrip_and_reference_reads_our_file(cuda_device, tmp_path, metric):
    """larger index, several read groups: write -> read gives identical search results; where the
    compiled reference is present it reads OUR file and agrees on lists and results"""
    import hnsw_ivf_b200 as b2
    from oracle import pyoracle as po
    d, nb, nlist, nprobe, k, nq = 48, 120000, 50, 6, 10, 200
    xb = clustered(nb, d, 4 * nlist, 3, normalize=metric == 0)
    xq = clustered(nq, d, 4 * nlist, 4, normalize=metric == 0)
    ids = np.arange(nb, dtype=np.int64) * 3 + 1
    idx = b2.IndexIVFFlat(d, nlist, metric)
    idx.set_centroids(sample_centroids(xb, nlist))
    idx.add_with_ids(xb, ids)
    idx.nprobe = nprobe
    out = str(tmp_path / "big.ivfflat")
    b2.write_index(idx, out)
    idx2 = b2.read_index(out)
    assert idx2.ntotal == nb
    D1, I1 = idx.search(xq, k)
    D2, I2 = idx2.search(xq, k)
    assert np.array_equal(I1, I2) and np.array_equal(D1, D2)
    if not po.ref_available():
        pytest.skip("oracle/_ref not built: reference-side read not checked")
    r = po.RefIndex.read_index(out)
    assert (r.d, r.nlist, r.nprobe, r.ntotal, r.metric) == (d, nlist, nprobe, nb, metric)
    for l in (0, 7, nlist - 1):
        cr, ir = r.get_list(l)
        cg, ig = idx.get_list(l)
        assert np.array_equal(ir, ig) and np.array_equal(cr, cg)
    _, Iq = idx.quantize(xq, nprobe)
    Dr, Ir = r.search_preassigned(xq, k, Iq)
    Dg, Ig = idx.search_preassigned(xq, k, Iq)
    assert_knn_parity(Dr, Ir, Dg, Ig, what="reference on our file")


def test_read_errors(cuda_device, tmp_path):
    """error behaviour of read_index: code -2 (FaissException) with a message, no partial handle"""
    import hnsw_ivf_b200 as b2
    with pytest.raises(b2.B200Error) as e:
        b2.read_index(str(tmp_path / "missing.ivfflat"))
    assert e.value.code == -2 and "could not open" in str(e.value)
    p = tmp_path / "flat.index"
    p.write_bytes(b"IxF2" + b"\0" * 64)  # an IndexFlat file: not this path's index type
    with pytest.raises(b2.B200Error) as e:
        b2.read_index(str(p))
    assert "not recognized" in str(e.value)
    good = open(os.path.join(GOLDEN_DIR, "lowlevel_l2.ivfflat"), "rb").read()
    p = tmp_path / "truncated.ivfflat"
    p.write_bytes(good[: len(good) // 2])
    with pytest.raises(b2.B200Error) as e:
        b2.read_index(str(p))
    assert "read error" in str(e.value)
