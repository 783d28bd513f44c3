"""GPU parity tests: the CUDA path, called through the C-ABI, against the CPU oracle on the same
seeded inputs. Shapes follow the reference's own tests for this path (tests/test_lowlevel_ivf.cpp:
d=32, nb=1000, nq=200, k=10, nlist=32, nprobe=4) plus the edge cases it exercises."""
import numpy as np
import pytest

from helpers import (assert_knn_parity, brute_force, clustered, compare_knn, recall_at_k,
                     sample_centroids)

pytestmark = pytest.mark.gpu

L2, IP = 1, 0


def _oracle(d, nlist, metric, centroids, xb, ids=None, assign=None):
    from oracle import pyoracle as po
    o = po.PortIndex(d, nlist, metric)
    o.set_centroids(centroids)
    if assign is not None:
        o.add_core(xb, ids, assign)
    else:
        o.add_with_ids(xb, ids)
    return o


def _lists_assignment(o, nb):
    """per-vector list number as the oracle placed them (ids are 0..nb-1)"""
    a = np.full(nb, -1, np.int64)
    for l in range(o.nlist):
        _, ids = o.get_list(l)
        a[ids] = l
    return a


def _build_pair(d, nb, nlist, metric, seed=1, normalize=False):
    """oracle index + device index holding IDENTICAL lists (device filled with the oracle's
    assignment through add_core, so that search parity is not blurred by assignment near-ties)"""
    import hnsw_ivf_b200 as b2
    xb = clustered(nb, d, 4 * nlist, seed, normalize=normalize)
    cent = sample_centroids(xb, nlist)
    o = _oracle(d, nlist, metric, cent, xb)
    g = b2.IndexIVFFlat(d, nlist, metric)
    g.set_centroids(cent)
    g.add_core(xb, None, _lists_assignment(o, nb))
    return xb, cent, o, g


@pytest.mark.parametrize("metric", [L2, IP])
def test_lowlevel_shape(cuda_device, metric):
    d, nb, nq, k, nlist, nprobe = 32, 1000, 200, 10, 32, 4
    xb, cent, o, g = _build_pair(d, nb, nlist, metric)
    xq = clustered(nq, d, 4 * nlist, 99)
    for l in range(nlist):
        co, io = o.get_list(l)
        cg, ig = g.get_list(l)
        assert np.array_equal(io, ig) and np.array_equal(co, cg), "list %d differs" % l
    Dq, Iq = o.quantize(xq, nprobe)
    Do, Io = o.search_preassigned(xq, k, Iq)
    g.nprobe = nprobe
    Dg, Ig = g.search_preassigned(xq, k, Iq)
    assert_knn_parity(Do, Io, Dg, Ig, what="preassigned")
    # end to end (device coarse quantizer)
    De, Ie = g.search(xq, k)
    r = compare_knn(Do, Io, De, Ie)
    assert r["exact_id_rate"] > 0.99, r
    # stats: ndis is the reference's counter (faiss/IndexIVF.cpp:624,721)
    g.reset_stats()
    o.stats.ndis = 0
    o.search_preassigned(xq, k, Iq)
    g.search_preassigned(xq, k, Iq)
    s = g.get_stats()
    assert s.ndis == o.stats.ndis and s.nq == nq


@pytest.mark.parametrize("metric", [L2, IP])
@pytest.mark.parametrize("d,nb,nlist,nprobe,k,nq", [
    (128, 50000, 256, 16, 10, 300),
    (96, 20000, 64, 8, 10, 64),
    (30, 5000, 16, 3, 5, 33),     # d not a multiple of 4
    (7, 3000, 8, 8, 1, 17),       # tiny d, nprobe = nlist, k = 1
    (64, 8000, 40, 5, 100, 20),   # k >= 100
    (960, 4000, 16, 4, 10, 12),   # wide rows (GIST shape)
])
def test_search_parity(cuda_device, metric, d, nb, nlist, nprobe, k, nq):
    xb, cent, o, g = _build_pair(d, nb, nlist, metric, normalize=(metric == IP))
    xq = clustered(nq, d, 4 * nlist, 4321, normalize=(metric == IP))
    Dq, Iq = o.quantize(xq, nprobe)
    Do, Io = o.search_preassigned(xq, k, Iq)
    g.nprobe = nprobe
    Dg, Ig = g.search_preassigned(xq, k, Iq)
    assert_knn_parity(Do, Io, Dg, Ig, what="d=%d" % d)


@pytest.mark.parametrize("metric", [L2, IP])
def test_quantizer_parity(cuda_device, metric):
    import hnsw_ivf_b200 as b2
    from oracle import pyoracle as po
    d, nlist = 64, 512
    cent = clustered(nlist, d, 64, 3)
    xq = clustered(777, d, 64, 8)
    g = b2.IndexIVFFlat(d, nlist, metric)
    g.set_centroids(cent)
    for nprobe in (1, 8, 100):
        Dr, Ir = po.port_knn(metric, xq, cent, nprobe)
        Dg, Ig = g.quantize(xq, nprobe)
        r = compare_knn(Dr, Ir, Dg, Ig, rtol=1e-4, atol=1e-4)
        assert r["unexplained"] == 0 and r["bad_dist"] == 0, (nprobe, r)
        assert r["exact_id_rate"] > 0.995, (nprobe, r)
    # fewer than 20 queries: the reference switches to the direct form (distances.cpp:817-821)
    Dr, Ir = po.port_knn(metric, xq[:7], cent, 8)
    Dg, Ig = g.quantize(xq[:7], 8)
    assert compare_knn(Dr, Ir, Dg, Ig)["unexplained"] == 0


def test_exhaustive_equals_brute_force(cuda_device):
    # nprobe = nlist must be exact kNN (tests/test_index.py:303-345, strong form)
    d, nb, nlist, k = 48, 6000, 24, 10
    xb, cent, o, g = _build_pair(d, nb, nlist, L2)
    xq = clustered(50, d, 4 * nlist, 11)
    g.nprobe = nlist
    D, I = g.search(xq, k)
    Dt, It = brute_force(L2, xq, xb, k)
    assert recall_at_k(I, It) == 1.0
    np.testing.assert_allclose(D, Dt, rtol=1e-4, atol=1e-5)


def test_padding_and_skips(cuda_device):
    import hnsw_ivf_b200 as b2
    d, nlist, k = 16, 8, 10
    xb = clustered(12, d, 8, 1)
    cent = clustered(nlist, d, 8, 2)
    assign = np.array([0, 0, 0, 1, 1, 5, 5, 5, 5, 5, 5, -1], np.int64)  # lists 2,3,4,6,7 empty
    for metric in (L2, IP):
        o = _oracle(d, nlist, metric, cent, xb, None, assign)
        g = b2.IndexIVFFlat(d, nlist, metric)
        g.set_centroids(cent)
        g.add_core(xb, None, assign)
        assert g.ntotal == 12 and o.ntotal == 12
        xq = clustered(5, d, 8, 3)
        keys = np.array([[0, 1, -1], [2, 3, 4], [5, -1, -1], [1, 5, 0], [7, 6, 2]], np.int64)
        g.nprobe = 3
        Do, Io = o.search_preassigned(xq, k, keys)
        Dg, Ig = g.search_preassigned(xq, k, keys)
        assert np.array_equal(Io, Ig)
        assert np.array_equal(Io == -1, Ig == -1)
        np.testing.assert_allclose(Dg, Do, rtol=1e-4)
        neutral = np.finfo(np.float32).max if metric == L2 else -np.finfo(np.float32).max
        assert np.all(Dg[Ig == -1] == neutral)
        # invalid key: the reference throws (faiss/IndexIVF.cpp:520-524)
        bad = keys.copy()
        bad[0, 0] = nlist
        with pytest.raises(b2.B200Error) as ei:
            g.search_preassigned(xq, k, bad)
        assert ei.value.code == -2
        with pytest.raises(b2.B200Error):
            g.search(xq, 0)  # FAISS_THROW_IF_NOT(k > 0)


def test_exact_ties_order(cuda_device):
    """duplicated vectors: equal distances must come out in cmp2 order (L2: ascending id,
    IP: descending id), faiss/utils/Heap.h:420-450"""
    import hnsw_ivf_b200 as b2
    d, nlist, k = 8, 2, 6
    base = clustered(3, d, 2, 5)
    xb = np.concatenate([base, base, base])  # ids 0..8, id i duplicates i%3
    cent = clustered(nlist, d, 2, 6)
    assign = np.zeros(9, np.int64)
    xq = base[:1].copy()
    for metric in (L2, IP):
        o = _oracle(d, nlist, metric, cent, xb, None, assign)
        g = b2.IndexIVFFlat(d, nlist, metric)
        g.set_centroids(cent)
        g.add_core(xb, None, assign)
        g.nprobe = 1
        keys = np.zeros((1, 1), np.int64)
        Do, Io = o.search_preassigned(xq, k, keys)
        Dg, Ig = g.search_preassigned(xq, k, keys)
        np.testing.assert_allclose(Dg, Do, rtol=1e-5, atol=1e-6)
        # groups of equal distance hold the same ids in the same order
        assert np.array_equal(Ig, Io), (metric, Ig, Io)


def test_device_pointers_and_ragged_batches(cuda_device):
    import torch
    d, nb, nlist, nprobe, k = 64, 20000, 64, 8, 10
    xb, cent, o, g = _build_pair(d, nb, nlist, L2)
    g.nprobe = nprobe
    for nq in (1, 7, 33, 257):
        xq = clustered(nq, d, 4 * nlist, 100 + nq)
        Dh, Ih = g.search(xq, k)
        xd = torch.from_numpy(xq).cuda()
        Dd, Id = g.search(xd, k)
        assert Dd.is_cuda and Id.is_cuda
        assert np.array_equal(Id.cpu().numpy(), Ih)
        assert np.array_equal(Dd.cpu().numpy(), Dh)
        Dq, Iq = o.quantize(xq, nprobe)
        Do, Io = o.search_preassigned(xq, k, Iq)
        Dg, Ig = g.search_preassigned(xq, k, Iq)
        assert_knn_parity(Do, Io, Dg, Ig, min_exact=0.99, what="nq=%d" % nq)


def test_add_with_ids_assignment(cuda_device):
    """device top-1 assignment == reference assignment except at coarse near-ties"""
    import hnsw_ivf_b200 as b2
    d, nb, nlist = 32, 20000, 64
    xb = clustered(nb, d, 4 * nlist, 21)
    cent = sample_centroids(xb, nlist)
    ids = np.arange(nb, dtype=np.int64) * 3 + 1
    for metric in (L2, IP):
        o = _oracle(d, nlist, metric, cent, xb, ids)
        g = b2.IndexIVFFlat(d, nlist, metric)
        g.set_centroids(cent)
        g.add_with_ids(xb[:12345], ids[:12345])  # two adds: appends keep insertion order
        g.add_with_ids(xb[12345:], ids[12345:])
        assert g.ntotal == nb
        same = 0
        for l in range(nlist):
            _, io = o.get_list(l)
            cg, ig = g.get_list(l)
            common = np.intersect1d(io, ig)
            same += common.size
            # order inside a list is insertion order on both sides
            assert np.array_equal(ig, np.sort(ig))
            assert np.array_equal(cg, xb[(ig - 1) // 3])
        assert same >= nb - 5, same


def test_sharding_by_list_range_equals_single(cuda_device):
    """tests/test_meta_index.py:128-176 / faiss/gpu/test/test_multi_gpu.py:54-96: shards + merge ==
    single index. Two handles on one GPU stand for two ranks."""
    import hnsw_ivf_b200 as b2
    from oracle import pyoracle as po
    d, nb, nlist, nprobe, k, nq = 32, 10000, 32, 6, 10, 100
    for metric in (L2, IP):
        xb, cent, o, g = _build_pair(d, nb, nlist, metric)
        xq = clustered(nq, d, 4 * nlist, 17)
        assign = _lists_assignment(o, nb)
        shards = []
        for r in range(2):
            s = b2.IndexIVFFlat(d, nlist, metric)
            s.set_centroids(cent)
            s.set_list_range(*b2.list_range_of_rank(nlist, r, 2))
            s.add_core(xb, None, assign)
            s.nprobe = nprobe
            shards.append(s)
        g.nprobe = nprobe
        D1, I1 = g.search(xq, k)
        parts = [s.search(xq, k) for s in shards]
        all_D = np.stack([p[0] for p in parts])
        all_I = np.stack([p[1] for p in parts])
        Dm, Im = b2.merge_knn_results(metric, all_D, all_I)
        assert np.array_equal(Im, I1) and np.array_equal(Dm, D1)
        Dp, Ip = po.port_merge_knn_results(metric, all_D, all_I)
        assert np.array_equal(Ip, Im) and np.array_equal(Dp, Dm)


# ------------------------------------------------------------------------------------------------
# the committed golden vectors of the reference (tests/golden/make_golden.py), through the C-ABI
# ------------------------------------------------------------------------------------------------
import glob  # noqa: E402
import os  # noqa: E402

_GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


@pytest.mark.parametrize("path", _GOLDEN, ids=[os.path.basename(p)[:-4] for p in _GOLDEN])
def test_golden_vectors(cuda_device, path):
    import hnsw_ivf_b200 as b2
    g = np.load(path)
    metric, nprobe, k, nlist = int(g["metric"]), int(g["nprobe"]), int(g["k"]), int(g["nlist"])
    ties = "ties" in path
    idx = b2.IndexIVFFlat(g["xb"].shape[1], nlist, metric)
    idx.set_centroids(g["centroids"])
    idx.add_core(g["xb"], g["ids"], g["assign"])  # the reference's own list assignment
    idx.nprobe = nprobe
    for l in range(nlist):
        want = np.nonzero(g["assign"] == l)[0]
        codes, ids = idx.get_list(l)
        assert np.array_equal(ids, g["ids"][want]) and np.array_equal(codes, g["xb"][want])
    # scan + select with the reference's probe table: ids equal outside exact/near ties
    Dp, Ip = idx.search_preassigned(g["xq"], k, g["Iq"])
    assert_knn_parity(g["Dp"], g["Ip"], Dp, Ip, min_exact=0.0 if ties else 0.999, what="golden")
    if ties:
        # groups of exactly equal distance hold the same ids in the reference's cmp2 order,
        # except where the group is cut by the k-th boundary (which members survive depends on
        # scan order there: the documented exception)
        same = (Ip == g["Ip"]).mean()
        assert same > 0.9, same
    idx.reset_stats()
    idx.search_preassigned(g["xq"], k, g["Iq"])
    assert idx.get_stats().ndis == int(g["ndis"])
    # coarse quantizer vs the reference's BLAS path
    Dq, Iq = idx.quantize(g["xq"], nprobe)
    r = compare_knn(g["Dq"], g["Iq"], Dq, Iq, rtol=1e-4, atol=1e-4)
    assert r["unexplained"] == 0 and r["bad_dist"] == 0, r
    # end to end
    D, I = idx.search(g["xq"], k)
    assert_knn_parity(g["D"], g["I"], D, I, min_exact=0.0 if ties else 0.99, what="golden e2e")
    # device assignment (add_with_ids) == the reference's, outside coarse near-ties
    idx2 = b2.IndexIVFFlat(g["xb"].shape[1], nlist, metric)
    idx2.set_centroids(g["centroids"])
    idx2.add_with_ids(g["xb"], g["ids"])
    diff = sum(abs(idx2.get_list_size(l) - int((g["assign"] == l).sum())) for l in range(nlist))
    assert diff <= (2 * g["xb"].shape[0]) // 100 if ties else diff <= 4, diff


@pytest.mark.parametrize("metric", [L2, IP])
def test_small_batch_graph_replay(cuda_device, metric):
    """small batches run as one captured CUDA graph from the third call of a shape on: replays must
    give what the eager first call gave, for new queries too, and a graph must not outlive an add()
    that moves the slab (the low-latency regime of BASELINE config 5)"""
    import hnsw_ivf_b200 as b2
    d, nb, nlist, nprobe, k = 64, 20000, 128, 8, 10
    xb, cent, o, g = _build_pair(d, nb, nlist, metric)
    g.nprobe = nprobe
    xq = clustered(64, d, 4 * nlist, 7)
    for bs in (1, 5, 32):
        outs = []
        for r in range(4):  # eager, capture + replay, replay, replay
            q = xq[r:r + bs]
            Dq, Iq = o.quantize(q, nprobe)
            Do, Io = o.search_preassigned(q, k, Iq)
            Dg, Ig = g.search_preassigned(q, k, Iq)
            assert_knn_parity(Do, Io, Dg, Ig, what="graph preassigned bs=%d call %d" % (bs, r))
            De, Ie = g.search(q, k)
            r2 = compare_knn(Do, Io, De, Ie)
            assert r2["exact_id_rate"] > 0.97, r2
            outs.append((De, Ie))
    # grow the index: the slab moves, graphs captured before must be dropped
    xb2 = clustered(3000, d, 4 * nlist, 11)
    ids2 = np.arange(nb, nb + 3000, dtype=np.int64)
    o.add_with_ids(xb2, ids2)
    a2 = np.empty(3000, np.int64)
    for l in range(nlist):
        _, ids = o.get_list(l)
        sel = ids[ids >= nb] - nb
        a2[sel] = l
    g.add_core(xb2, ids2, a2)
    for r in range(3):
        q = xq[r:r + 5]
        Dq, Iq = o.quantize(q, nprobe)
        Do, Io = o.search_preassigned(q, k, Iq)
        Dg, Ig = g.search_preassigned(q, k, Iq)
        assert_knn_parity(Do, Io, Dg, Ig, what="graph after add, call %d" % r)


@pytest.mark.parametrize("metric", [L2, IP])
@pytest.mark.parametrize("nshard,k", [(2, 10), (3, 1), (8, 10), (8, 64), (5, 100), (40, 16)])
def test_merge_knn_results_tables(cuda_device, metric, nshard, k):
    """merge_knn_results (faiss/utils/Heap.cpp:158-232) on random sorted per-shard tables with -1
    padding and exact ties: both device kernels (rank merge for nshard*k <= 512, head walk above)
    against the restatement"""
    from oracle import pyoracle as po
    import hnsw_ivf_b200 as b2
    rng = np.random.default_rng(nshard * 100 + k)
    n = 257
    all_D = np.empty((nshard, n, k), np.float32)
    all_I = np.empty((nshard, n, k), np.int64)
    for s in range(nshard):
        v = rng.integers(0, 40 * k, size=(n, k)).astype(np.float32) / 4.0  # coarse grid: many ties
        v.sort(axis=1)
        if metric == IP:
            v = -v
        ids = rng.permutation(n * k * nshard)[: n * k].reshape(n, k).astype(np.int64) * nshard + s
        nvalid = rng.integers(0, k + 1, size=n)
        nvalid[: n // 2] = k
        for q in range(n):
            ids[q, nvalid[q]:] = -1
            v[q, nvalid[q]:] = np.finfo(np.float32).max if metric == L2 else -np.finfo(np.float32).max
        all_D[s], all_I[s] = v, ids
    Dg, Ig = b2.merge_knn_results(metric, all_D, all_I)
    Dp, Ip = po.port_merge_knn_results(metric, all_D, all_I)
    assert np.array_equal(Dg, Dp)
    assert np.array_equal(Ig < 0, Ip < 0)
    for q in range(n):
        if np.array_equal(Ig[q], Ip[q]):
            continue
        # ids may differ only inside runs of equal distance (ties: by shard there, by id here),
        # and then as sets unless the run is cut by the k-th slot
        for val in np.unique(Dg[q][Ig[q] >= 0]):
            m = Dg[q] == val
            if m[-1]:
                continue
            assert set(Ig[q][m].tolist()) == set(Ip[q][m].tolist()), (q, val)
        # every id we return is one of the inputs with that distance
        for j in range(k):
            if Ig[q, j] >= 0:
                s = int(Ig[q, j] % nshard)
                pos = np.nonzero(all_I[s, q] == Ig[q, j])[0]
                assert pos.size == 1 and all_D[s, q, pos[0]] == Dg[q, j]
