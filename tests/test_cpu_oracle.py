"""CPU tests (no GPU): the oracle against the reference's golden vectors and against the identities
the reference's own tests pin for this path. The oracle here is the plain-C
restatement oracle/ivfflat_oracle.c; where oracle/_ref (the compiled, unmodified reference) is
present it is checked too, which is what pins the restatement."""
import glob
import os

import numpy as np
import pytest

from helpers import (assert_knn_parity, brute_force, clustered, compare_knn, recall_at_k,
                     sample_centroids)
from oracle import pyoracle as po

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))
L2, IP = 1, 0


def _port_from_golden(g):
    d = g["xb"].shape[1]
    o = po.PortIndex(d, int(g["nlist"]), int(g["metric"]))
    o.set_centroids(g["centroids"])
    return o


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_port_matches_reference_golden(path):
    """every stage of the path, port vs the outputs the reference produced for the same inputs"""
    g = np.load(path)
    metric, nprobe, k = int(g["metric"]), int(g["nprobe"]), int(g["k"])
    o = _port_from_golden(g)
    o.add_with_ids(g["xb"], g["ids"])
    # (i) list membership and order (faiss/IndexIVFFlat.cpp:46-91: insertion order inside a list)
    pos = {int(v): i for i, v in enumerate(g["ids"])}
    for l in range(int(g["nlist"])):
        codes, ids = o.get_list(l)
        want = np.nonzero(g["assign"] == l)[0]
        got = np.array([pos[int(v)] for v in ids], dtype=np.int64)
        if "ties" not in path:  # duplicated vectors may tie between two centroids
            assert np.array_equal(got, want), "list %d membership/order" % l
            assert np.array_equal(codes, g["xb"][want])
    # (ii) probe table
    Dq, Iq = o.quantize(g["xq"], nprobe)
    r = compare_knn(g["Dq"], g["Iq"], Dq, Iq, rtol=1e-4, atol=1e-5)
    assert r["unexplained"] == 0 and r["bad_dist"] == 0, r
    # (iii) preassigned search with the REFERENCE's probes: ids position-wise equal outside ties
    Dp, Ip = o.search_preassigned(g["xq"], k, g["Iq"])
    r = assert_knn_parity(g["Dp"], g["Ip"], Dp, Ip, min_exact=0.0 if "ties" in path else 0.999,
                          what="preassigned")
    if "ties" in path:
        # exact ties: the reference's cmp2 order (faiss/utils/Heap.h:420-450) must be reproduced
        assert np.array_equal(Ip, g["Ip"]), "tie order differs from the reference"
    assert o.stats.ndis == int(g["ndis"])
    # (iv) end to end
    D, I = o.search(g["xq"], k, nprobe)
    assert_knn_parity(g["D"], g["I"], D, I, min_exact=0.0 if "ties" in path else 0.995,
                      what="search")


@pytest.mark.parametrize("path", GOLDEN[:3], ids=[os.path.basename(p)[:-4] for p in GOLDEN[:3]])
def test_golden_decomposition_identity(path):
    """tests/test_lowlevel_ivf.cpp:76-216: index->search == quantizer->search + per-probe scan +
    reorder; checked on the golden outputs themselves (D/I vs Dp/Ip) and on the port"""
    g = np.load(path)
    assert np.array_equal(g["I"], g["Ip"]) and np.array_equal(g["D"], g["Dp"])


@pytest.mark.skipif(not po.ref_available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("metric", [L2, IP])
def test_port_matches_compiled_reference(metric):
    d, nb, nq, nlist, nprobe, k = 48, 6000, 150, 40, 6, 10
    xb = clustered(nb, d, 4 * nlist, 3, normalize=(metric == IP))
    xq = clustered(nq, d, 4 * nlist, 4, normalize=(metric == IP))
    cent = sample_centroids(xb, nlist)
    o = po.PortIndex(d, nlist, metric)
    r = po.RefIndex(d, nlist, metric)
    for x in (o, r):
        x.set_centroids(cent)
        x.add_with_ids(xb)
    for l in range(nlist):
        a, b = o.get_list(l), r.get_list(l)
        assert np.array_equal(a[1], b[1]) and np.array_equal(a[0], b[0])
    Dq, Iq = r.quantize(xq, nprobe)
    Do, Io = o.search_preassigned(xq, k, Iq)
    Dr, Ir = r.search_preassigned(xq, k, Iq)
    assert_knn_parity(Dr, Ir, Do, Io, min_exact=0.999)
    # parallel_mode invariance (tests/test_index_accuracy.py:54-60)
    r.set_parallel_mode(1)
    D1, I1 = r.search(xq, k, nprobe)
    r.set_parallel_mode(0)
    D0, I0 = r.search(xq, k, nprobe)
    assert np.array_equal(D0, D1) and (I0 != I1).sum() <= I0.size // 5000 + 1


@pytest.mark.parametrize("metric", [L2, IP])
def test_probe_partition_merge_identity(metric):
    """tests/test_lowlevel_ivf.cpp:398-549: probes split over 3 workers, private heaps, merged ==
    the single search. This is the property list-range sharding relies on."""
    d, nb, nq, nlist, nprobe, k = 32, 3000, 60, 24, 6, 10
    xb = clustered(nb, d, 4 * nlist, 5)
    xq = clustered(nq, d, 4 * nlist, 6)
    o = po.PortIndex(d, nlist, metric)
    o.set_centroids(sample_centroids(xb, nlist))
    o.add_with_ids(xb)
    Dq, Iq = o.quantize(xq, nprobe)
    D, I = o.search_preassigned(xq, k, Iq)
    parts = []
    for w in range(3):
        keys = np.full_like(Iq, -1)
        keys[:, w::3] = Iq[:, w::3]
        parts.append(o.search_preassigned(xq, k, keys))
    Dm, Im = po.port_merge_knn_results(metric, np.stack([p[0] for p in parts]),
                                       np.stack([p[1] for p in parts]))
    assert np.array_equal(Im, I) and np.array_equal(Dm, D)


def test_exhaustive_probe_is_brute_force():
    """tests/test_index.py:303-345 (strong form): nprobe = nlist is exact kNN"""
    d, nb, nlist, k = 24, 4000, 16, 10
    xb = clustered(nb, d, 64, 7)
    xq = clustered(40, d, 64, 8)
    o = po.PortIndex(d, nlist, L2)
    o.set_centroids(sample_centroids(xb, nlist))
    o.add_with_ids(xb)
    D, I = o.search(xq, k, nlist)
    Dt, It = brute_force(L2, xq, xb, k)
    assert recall_at_k(I, It) == 1.0
    np.testing.assert_allclose(D, Dt, rtol=1e-4, atol=1e-5)


def test_heap_semantics():
    """strict compare on insert, cmp2 order on reorder, neutral padding
    (faiss/utils/Heap.h:111-150,420-450; faiss/utils/ordered_key_value.h:41-84)"""
    vals = np.array([3, 1, 2, 1, 3, 1, 0.5], np.float32)
    ids = np.array([10, 11, 12, 5, 9, 20, 7], np.int64)
    v, i = po.port_heap_select(L2, 4, vals, ids)
    assert v.tolist() == [0.5, 1, 1, 1] and i.tolist() == [7, 5, 11, 20]
    v, i = po.port_heap_select(IP, 3, vals, ids)
    assert v.tolist() == [3, 3, 2] and i.tolist() == [10, 9, 12]  # ties: descending id
    v, i = po.port_heap_select(L2, 10, vals[:2], ids[:2])
    assert i.tolist() == [11, 10] + [-1] * 8 and np.all(v[2:] == np.finfo(np.float32).max)
    v, i = po.port_heap_select(IP, 4, vals[:1], ids[:1])
    assert i.tolist() == [10, -1, -1, -1] and np.all(v[1:] == -np.finfo(np.float32).max)
    if po.ref_available():
        rng = np.random.default_rng(0)
        for metric in (L2, IP):
            vv = rng.integers(0, 6, 200).astype(np.float32)  # many ties
            ii = rng.permutation(200).astype(np.int64)
            a = po.port_heap_select(metric, 17, vv, ii)
            b = po.ref_heap_select(metric, 17, vv, ii)
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


@pytest.mark.parametrize("metric", [L2, IP])
def test_merge_knn_results(metric):
    """faiss/utils/Heap.cpp:158-232: sorted per-shard tables, -1 skipped"""
    rng = np.random.default_rng(1)
    nshard, n, k = 3, 50, 8
    D = rng.random((nshard, n, k)).astype(np.float32)
    D.sort(axis=2)
    if metric == IP:
        D = D[:, :, ::-1].copy()
    I = rng.permutation(nshard * n * k).reshape(nshard, n, k).astype(np.int64)
    I[1, :, 5:] = -1
    D[1, :, 5:] = np.finfo(np.float32).max if metric == L2 else -np.finfo(np.float32).max
    Dm, Im = po.port_merge_knn_results(metric, D, I)
    for q in range(n):
        cand = [(D[s, q, j], I[s, q, j]) for s in range(nshard) for j in range(k)
                if I[s, q, j] >= 0]
        cand.sort(key=lambda t: (t[0] if metric == L2 else -t[0]))
        assert [c[1] for c in cand[:k]] == Im[q].tolist()
    if po.ref_available():
        Dr, Ir = po.ref_merge_knn_results(metric, D, I)
        assert np.array_equal(Ir, Im) and np.array_equal(Dr, Dm)


def test_skips_and_padding():
    """key < 0 and empty lists are skipped, short results padded (faiss/IndexIVF.cpp:516-529,
    faiss/utils/Heap.h:445-448)"""
    d, nlist, k = 8, 6, 5
    xb = clustered(7, d, 4, 1)
    o = po.PortIndex(d, nlist, L2)
    o.set_centroids(clustered(nlist, d, 4, 2))
    o.add_core(xb, None, np.array([0, 0, 3, 3, 3, -1, 5], np.int64))
    assert o.ntotal == 7 and [o.list_size(l) for l in range(nlist)] == [2, 0, 0, 3, 0, 1]
    keys = np.array([[0, -1], [1, 2], [3, 5]], np.int64)
    D, I = o.search_preassigned(clustered(3, d, 4, 3), k, keys)
    assert (I[0] >= 0).sum() == 2 and (I[1] >= 0).sum() == 0 and (I[2] >= 0).sum() == 4
    assert np.all(D[I < 0] == np.finfo(np.float32).max)


# ---- the reference's index file format ------------------------------------
GOLDEN_DIR = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("name", ["lowlevel_l2", "lowlevel_ip"])
def test_index_file_restatement_matches_reference_file(name):
    """oracle.parse_index_file (the restated "IwFl" layout, faiss/impl/index_write.cpp:635-638)
    against a file the unmodified reference wrote for a golden case"""
    g = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    f = po.parse_index_file(os.path.join(GOLDEN_DIR, name + ".ivfflat"))
    assert (f["d"], f["nlist"], f["nprobe"], f["metric"]) == (
        g["xb"].shape[1], int(g["nlist"]), int(g["nprobe"]), int(g["metric"]))
    assert f["ntotal"] == g["xb"].shape[0] and f["is_trained"]
    assert f["quantizer"]["ntotal"] == f["nlist"] and f["quantizer"]["metric"] == f["metric"]
    assert f["quantizer"]["fourcc"] == po._fourcc("IxF2" if f["metric"] == 1 else "IxFI")
    assert np.array_equal(f["centroids"], g["centroids"])
    assert f["direct_map_type"] == 0
    for l in range(f["nlist"]):
        want = np.nonzero(g["assign"] == l)[0]
        codes, ids = f["lists"][l]
        assert np.array_equal(ids, g["ids"][want]) and np.array_equal(codes, g["xb"][want])


def test_index_file_sparse_size_table():
    """more than half of the lists empty: "sprs" table (faiss/impl/index_write.cpp:268-279)"""
    g = np.load(os.path.join(GOLDEN_DIR, "sparse_l2.inputs"))
    f = po.parse_index_file(os.path.join(GOLDEN_DIR, "sparse_l2.ivfflat"))
    assert f["nlist"] == 40 and f["nprobe"] == 5 and f["ntotal"] == 300
    for l in range(40):
        want = np.nonzero(g["assign"] == l)[0]
        codes, ids = f["lists"][l]
        assert np.array_equal(ids, g["ids"][want]) and np.array_equal(codes, g["xb"][want])


@pytest.mark.skipif(not po.ref_available(), reason="oracle/_ref not built")
def test_reference_reads_its_own_golden_file():
    """the committed file is a live reference index: read_index + search == the golden outputs"""
    g = np.load(os.path.join(GOLDEN_DIR, "lowlevel_l2.npz"))
    r = po.RefIndex.read_index(os.path.join(GOLDEN_DIR, "lowlevel_l2.ivfflat"))
    assert (r.d, r.nlist, r.nprobe, r.ntotal) == (32, 32, 4, 1000)
    po.ref_set_threads(1)
    D, I = r.search(g["xq"], int(g["k"]), int(g["nprobe"]))
    assert np.array_equal(I, g["I"]) and np.array_equal(D, g["D"])
