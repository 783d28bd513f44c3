"""GPU tests for the callers' side of the path, through the C-ABI:
SearchParametersIVF (IDSelector family on device, max_codes), store_pairs, range search, the
fork's search_stats, and k-means training -- against golden vectors of the UNMODIFIED reference
(tests/golden/params) and against the C restatement on fresh, larger seeded inputs."""
import glob
import json
import os

import numpy as np
import pytest

from helpers import (assert_knn_parity, clustered, compare_knn, compare_range, gpu_selector,
                     recall_at_k, resolve_spec, sample_centroids)

pytestmark = pytest.mark.gpu
PARAMS = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "params", "*.npz")))
L2, IP = 1, 0


def _gpu_from_golden(g):
    import hnsw_ivf_b200 as b2
    metric, d, nb, nq, nlist, nprobe, k = [int(v) for v in g["meta"]]
    ix = b2.IndexIVFFlat(d, nlist, metric)
    ix.set_centroids(g["centroids"])
    ix.add_core(g["xb"], g["ids"], g["assign"])
    ix.nprobe = nprobe
    return ix, (metric, d, nb, nq, nlist, nprobe, k)


@pytest.mark.parametrize("path", PARAMS, ids=[os.path.basename(p) for p in PARAMS])
def test_search_params_golden(cuda_device, path):
    import hnsw_ivf_b200 as b2
    g = dict(np.load(path))
    ix, (metric, d, nb, nq, nlist, nprobe, k) = _gpu_from_golden(g)
    for i, c in enumerate(json.loads(str(g["cases"]))):
        sel = gpu_selector(resolve_spec(c["sel"], g))
        p = b2.SearchParametersIVF(sel=sel, nprobe=nprobe, max_codes=c["max_codes"])
        ix.reset_stats()
        D, I = ix.search_preassigned(g["xq"], k, g["Iq"], store_pairs=c["store_pairs"], params=p)
        assert_knn_parity(g["D_%d" % i], g["I_%d" % i], D, I, what="case %d %s" % (i, c))
        assert ix.get_stats().ndis == int(g["ndis_%d" % i][0]), c


@pytest.mark.parametrize("path", PARAMS, ids=[os.path.basename(p) for p in PARAMS])
def test_range_search_golden(cuda_device, path):
    import hnsw_ivf_b200 as b2
    g = dict(np.load(path))
    ix, (metric, d, nb, nq, nlist, nprobe, k) = _gpu_from_golden(g)
    for i, c in enumerate(json.loads(str(g["range_cases"]))):
        sel = gpu_selector(resolve_spec(c["sel"], g))
        p = b2.SearchParametersIVF(sel=sel, nprobe=nprobe) if sel is not None else None
        l, D, I = ix.range_search_preassigned(g["xq"], c["radius"], g["Iq"],
                                              store_pairs=c["store_pairs"], params=p)
        nbd = compare_range(g["rl_%d" % i], g["rD_%d" % i], g["rI_%d" % i], l, D, I, c["radius"],
                            metric)
        assert nbd <= 2, (c, nbd)
        if nbd == 0:  # the reference's order too: probe order, then list order
            assert np.array_equal(l, g["rl_%d" % i]) and np.array_equal(I, g["rI_%d" % i]), c


def _pair(d, nb, nlist, metric, seed=51):
    import hnsw_ivf_b200 as b2
    from oracle import pyoracle as po
    xb = clustered(nb, d, 4 * nlist, seed, normalize=metric == IP)
    cent = sample_centroids(xb, nlist)
    ids = np.random.default_rng(seed).permutation(3 * nb)[:nb].astype(np.int64)
    o = po.PortIndex(d, nlist, metric)
    o.set_centroids(cent)
    o.add_with_ids(xb, ids)
    pos = {int(v): i for i, v in enumerate(ids)}
    a = np.full(nb, -1, np.int64)
    for l in range(nlist):
        a[[pos[int(v)] for v in o.get_list(l)[1]]] = l
    g = b2.IndexIVFFlat(d, nlist, metric)
    g.set_centroids(cent)
    g.add_core(xb, ids, a)
    return xb, ids, o, g


@pytest.mark.parametrize("metric", [L2, IP])
@pytest.mark.parametrize("d,nb,nlist,nprobe,k,nq", [
    (128, 60000, 128, 16, 10, 700),   # many queries per list: would be the tensor-core regime
    (30, 8000, 32, 5, 40, 50),        # d % 4 != 0, k > 32
    (64, 20000, 512, 32, 10, 9),      # few queries, lists hardly shared
])
def test_search_params_vs_port(cuda_device, metric, d, nb, nlist, nprobe, k, nq):
    import hnsw_ivf_b200 as b2
    xb, ids, o, g = _pair(d, nb, nlist, metric)
    xq = clustered(nq, d, 4 * nlist, 52, normalize=metric == IP)
    Dq, Iq = o.quantize(xq, nprobe)
    rng = np.random.default_rng(53)
    bitmap = rng.integers(0, 256, size=(3 * nb) // 8).astype(np.uint8)
    specs = [("range", nb // 2, 2 * nb), ("batch", ids[::5]), ("bitmap", bitmap),
             ("and", ("not", ("range", 0, nb)), ("bitmap", bitmap)), None]
    mean_list = nb // nlist
    for sp in specs:
        for mc in (0, mean_list // 2, 3 * mean_list + 7):
            if sp is None and mc == 0:
                continue
            p = b2.SearchParametersIVF(sel=gpu_selector(sp), nprobe=nprobe, max_codes=mc)
            o.stats.ndis = 0
            Do, Io = o.search_preassigned_ex(xq, k, Iq, max_codes=mc, sel=sp)
            g.reset_stats()
            tc0 = b2.tensor_core_launches()
            Dg, Ig = g.search_preassigned(xq, k, Iq, params=p)
            # filtered searches keep the list-major reuse: with many queries per list they run on
            # the rows-major tensor-core kernel (mask / row-limit arms), else on the exact kernel
            assert (b2.tensor_core_launches() > tc0) == (nq * nprobe > 2 * nlist)
            assert_knn_parity(Do, Io, Dg, Ig, what="%s mc=%d" % (sp and sp[0], mc))
            assert g.get_stats().ndis == o.stats.ndis
    # max_codes as the index field, end to end with the device quantizer and host/device buffers
    g.nprobe = nprobe
    g.max_codes = 2 * mean_list
    Do, Io = o.search_preassigned_ex(xq, k, Iq, max_codes=2 * mean_list)
    Dg, Ig = g.search(xq, k)
    r = compare_knn(Do, Io, Dg, Ig)
    assert r["exact_id_rate"] > 0.97, r
    g.max_codes = 0
    # store_pairs on every scan path (tensor cores included): labels are list << 32 | offset
    Do, Io = o.search_preassigned_ex(xq, k, Iq, store_pairs=True)
    Dg, Ig = g.search_preassigned(xq, k, Iq, store_pairs=True)
    assert_knn_parity(Do, Io, Dg, Ig, what="store_pairs")
    hit = Ig[Ig >= 0]
    lists, offs = hit >> 32, hit & 0xFFFFFFFF
    assert lists.max() < nlist and all(offs[i] < o.list_size(int(lists[i])) for i in range(0, hit.size, 37))


def test_params_errors(cuda_device):
    import hnsw_ivf_b200 as b2
    xb, ids, o, g = _pair(16, 2000, 8, L2)
    xq = clustered(5, 16, 32, 3)
    Dq, Iq = o.quantize(xq, 2)
    sel = b2.IDSelectorRange(0, 10)
    p = b2.SearchParametersIVF(sel=sel, nprobe=2)
    with pytest.raises(b2.B200Error) as e:  # faiss/IndexIVF.cpp:428-430
        g.search_preassigned(xq, 3, Iq, store_pairs=True, params=p)
    assert e.value.code == -2 and "store_pairs" in str(e.value)
    with pytest.raises(b2.B200Error):       # nprobe > 0, faiss/IndexIVF.cpp:414
        g.search(xq, 3, params=b2.SearchParametersIVF(nprobe=0))
    assert sel.is_member(3) and not sel.is_member(10)
    # a selector that keeps nothing: every slot is the padding (-1 / neutral distance)
    D, I = g.search(xq, 3, params=b2.SearchParametersIVF(sel=b2.IDSelectorRange(-5, -1), nprobe=8))
    assert np.all(I == -1) and np.all(D == np.finfo(np.float32).max)
    # params nprobe overrides the handle's
    g.nprobe = 1
    D8, I8 = g.search(xq, 3, params=b2.SearchParametersIVF(nprobe=8))
    g.nprobe = 8
    D8b, I8b = g.search(xq, 3)
    assert np.array_equal(I8, I8b)


@pytest.mark.parametrize("metric", [L2, IP])
def test_range_search_vs_port(cuda_device, metric):
    import hnsw_ivf_b200 as b2
    d, nb, nlist, nprobe, nq = 48, 30000, 64, 8, 300
    xb, ids, o, g = _pair(d, nb, nlist, metric)
    xq = clustered(nq, d, 4 * nlist, 61, normalize=metric == IP)
    xq[:3] = 1e3  # far from everything (L2): queries with no result at all
    Dq, Iq = o.quantize(xq, nprobe)
    D10, _ = o.search_preassigned(xq, 10, Iq)
    radius = float(np.median(D10[3:, 7]))
    g.nprobe = nprobe
    for sp in (None, ("range", 0, nb), ("not", ("batch", ids[::2]))):
        lo, Do, Io = o.range_search_preassigned(xq, radius, Iq, sel=sp)
        p = b2.SearchParametersIVF(sel=gpu_selector(sp), nprobe=nprobe) if sp else None
        lg, Dg, Ig = g.range_search_preassigned(xq, radius, Iq, params=p)
        nbd = compare_range(lo, Do, Io, lg, Dg, Ig, radius, metric)
        assert nbd <= 4 and int(lo[-1]) > nq, (sp and sp[0], nbd, int(lo[-1]))
        if nbd == 0:
            assert np.array_equal(lo, lg) and np.array_equal(Io, Ig)
    # end to end (device quantizer) and device-resident queries
    import torch
    lg2, Dg2, Ig2 = g.range_search(torch.from_numpy(xq).cuda(), radius)
    lo, Do, Io = o.range_search_preassigned(xq, radius, Iq)
    assert abs(int(lg2[-1]) - int(lo[-1])) <= 0.02 * int(lo[-1]) + 4
    # nothing inside the radius
    empty_r = -1.0 if metric == L2 else 1e9
    l0, D0, I0 = g.range_search(xq, empty_r)
    assert int(l0[-1]) == 0 and D0.size == 0 and np.all(l0 == 0)


def test_search_stats(cuda_device):
    """the fork's search_stats: same results as search, one QueryLatencyStats per query whose
    amortised figures add up to the batch's device time"""
    xb, ids, o, g = _pair(64, 40000, 64, L2)
    xq = clustered(500, 64, 256, 71)
    g.nprobe = 8
    D, I = g.search(xq, 10)
    Ds, Is, st = g.search_stats(xq, 10)
    assert np.array_equal(I, Is) and np.array_equal(D, Ds)
    t = g.last_timings()
    assert np.all(st["total_us"] > 0) and np.all(st["quantization_us"] > 0)
    np.testing.assert_allclose(st["quantization_us"].sum(), t["coarse"] * 1e3, rtol=1e-6)
    np.testing.assert_allclose(st["list_scan_us"].sum(),
                               (t["plan"] + t["scan"] + t["select"]) * 1e3, rtol=1e-6)
    np.testing.assert_allclose(st["total_us"], st["quantization_us"] + st["list_scan_us"])
    # a query that scans more rows is charged more
    Dq, Iq = g.quantize(xq, 8)
    nscan = np.array([sum(g.get_list_size(int(l)) for l in row) for row in Iq])
    assert np.argmax(st["list_scan_us"]) == np.argmax(nscan)


@pytest.mark.parametrize("path", PARAMS, ids=[os.path.basename(p) for p in PARAMS])
def test_train_golden(cuda_device, path):
    """same permutations, same summation order: the device k-means reproduces the reference's
    centroids unless a top-1 assignment flips on a near-tie (the reference assigns through the
    ||x||^2 + ||c||^2 - 2xc form, the device through exact differences), so most centroids are
    equal to rounding and all are close"""
    import hnsw_ivf_b200 as b2
    g = dict(np.load(path))
    metric, d, nb, nq, nlist, nprobe, k = [int(v) for v in g["meta"]]
    ix = b2.IndexIVFFlat(d, nlist, metric)
    assert not ix.is_trained
    ix.train(g["xt"])
    assert ix.is_trained and ix.ntotal == 0
    c = ix.get_centroids()
    err = np.abs(c - g["trained"]).max(axis=1)
    assert np.mean(err < 1e-5) >= 0.75 and err.max() < 0.2, (np.mean(err < 1e-5), err.max())
    ix.train(g["xt"][:100])  # trained already: a no-op (faiss/IndexIVF.cpp:72-74)
    assert np.array_equal(ix.get_centroids(), c)
    with pytest.raises(b2.B200Error):
        b2.IndexIVFFlat(d, nlist, metric).train(g["xt"][:nlist - 1])  # fewer points than clusters


@pytest.mark.parametrize("metric", [L2, IP])
def test_train_objective_and_recall(cuda_device, metric):
    """larger than the subsampling threshold, device-resident input: the k-means objective is
    within 1 % of the restatement's (same picks, same update), and train -> add -> search reaches
    the recall the restatement reaches with its own centroids"""
    import torch
    import hnsw_ivf_b200 as b2
    from oracle import pyoracle as po
    from helpers import brute_force
    d, nlist, nb, nq, k, nprobe = 32, 64, 40000, 200, 10, 8
    xb = clustered(nb, d, 4 * nlist, 81, normalize=metric == IP)
    xq = clustered(nq, d, 4 * nlist, 82, normalize=metric == IP)
    o = po.PortIndex(d, nlist, metric)
    o.train(xb)  # 40000 > 64 * 256: subsampled to 16384 points
    co = o.get_centroids()
    ix = b2.IndexIVFFlat(d, nlist, metric)
    ix.train(torch.from_numpy(xb).cuda())
    cg = ix.get_centroids()

    def objective(c):
        if metric == L2:
            x64, c64 = xb.astype(np.float64), c.astype(np.float64)
            dis = (x64 ** 2).sum(1)[:, None] + (c64 ** 2).sum(1)[None, :] - 2 * x64 @ c64.T
            return float(dis.min(1).sum())
        return float((xb @ c.T).max(1).sum())
    fo, fg = objective(co), objective(cg)
    assert abs(fo - fg) <= 0.01 * abs(fo), (fo, fg)
    assert np.mean(np.abs(co - cg).max(axis=1) < 1e-4) > 0.5
    if metric == IP:
        np.testing.assert_allclose(np.linalg.norm(cg, axis=1), 1.0, atol=1e-5)  # spherical
    ix.add(xb)
    ix.nprobe = nprobe
    D, I = ix.search(xq, k)
    o.add_with_ids(xb)
    Do, Io = o.search(xq, k, nprobe)
    _, gt = brute_force(metric, xq, xb, k)
    assert recall_at_k(I, gt) >= recall_at_k(Io, gt) - 0.02


def test_more_queries_than_one_device_batch(cuda_device):
    """n > 32768 queries: the host loop over device batches (range-search result offsets carried
    across batches, per-query stats filled batch by batch, selector mask evaluated once)"""
    import hnsw_ivf_b200 as b2
    d, nb, nlist, nprobe, k, nq = 16, 20000, 32, 4, 5, 40000
    xb, ids, o, g = _pair(d, nb, nlist, L2, seed=91)
    xq = clustered(nq, d, 4 * nlist, 92)
    Dq, Iq = o.quantize(xq, nprobe)
    g.nprobe = nprobe
    sp = ("range", 0, 2 * nb)
    p = b2.SearchParametersIVF(sel=gpu_selector(sp), nprobe=nprobe, max_codes=1500)
    Do, Io = o.search_preassigned_ex(xq, k, Iq, max_codes=1500, sel=sp)
    Dg, Ig = g.search_preassigned(xq, k, Iq, params=p)
    assert_knn_parity(Do, Io, Dg, Ig, what="two batches, params")
    radius = float(np.median(Do[:, 2]))
    lo, Dr, Ir = o.range_search_preassigned(xq, radius, Iq)
    lg, Dgr, Igr = g.range_search_preassigned(xq, radius, Iq)
    assert int(lo[-1]) > nq
    nbd = compare_range(lo, Dr, Ir, lg, Dgr, Igr, radius, L2)
    assert nbd <= 8, nbd
    if nbd == 0:
        assert np.array_equal(lo, lg) and np.array_equal(Ir, Igr)
    Ds, Is, st = g.search_stats(xq, k)
    D2, I2 = g.search(xq, k)
    assert np.array_equal(Is, I2) and st.shape == (nq,) and np.all(st["total_us"] > 0)
