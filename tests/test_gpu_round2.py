"""GPU tests of what round 2 added: the k-select pivot rule for k above the number of partial lists,
the tensor-core path for k / nprobe up to 248 (the reference's reservoir range,
faiss/impl/ResultHandler.h:369-477), list reservations with in-place append, asynchronous
all-device searches (deferred invalid-key report), the near-tie behaviour of the 3xTF32 selection,
and the sharded index over NCCL against the single index (faiss/gpu/test/test_multi_gpu.py:54-96).
Everything goes through the C-ABI (ctypes)."""
import os
import socket
import sys

import numpy as np
import pytest

from helpers import assert_knn_parity, clustered, compare_knn, sample_centroids

pytestmark = pytest.mark.gpu
L2, IP = 1, 0
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _oracle_and_assign(d, nb, nlist, metric, xb, cent):
    from oracle import pyoracle as po
    o = po.PortIndex(d, nlist, metric)
    o.set_centroids(cent)
    o.add_with_ids(xb)
    assign = np.full(nb, -1, np.int64)
    for l in range(nlist):
        assign[o.get_list(l)[1]] = l
    return o, assign


def _gpu(d, nlist, metric, cent, xb, assign, **opts):
    import hnsw_ivf_b200 as b2
    g = b2.IndexIVFFlat(d, nlist, metric)
    for name, v in opts.items():
        g.set_option(name, v)
    g.set_centroids(cent)
    g.add_core(xb, None, assign)
    return g


# ---- ADVICE (high): k above the number of partial lists -----------------------------------------
@pytest.mark.parametrize("metric", [L2, IP])
@pytest.mark.parametrize("nprobe,k,nq,use_tc", [
    (32, 100, 700, 0),   # L = nprobe * S is a power of two >= 32 and k > L: the old head bound
    (64, 100, 700, 0),   #   kept fewer than k keys and returned -1 padding
    (32, 100, 700, 1),   # the same through the tensor-core partial lists (KP = 128)
    (16, 200, 40, 0),    # few queries, wide merge (one CTA per query), k > nprobe * S
    (8, 64, 300, 1),
])
def test_select_k_above_partial_lists(cuda_device, metric, nprobe, k, nq, use_tc):
    d, nb, nlist = 32, 60000, 64
    xb = clustered(nb, d, 4 * nlist, 11, normalize=(metric == IP))
    cent = sample_centroids(xb, nlist)
    o, assign = _oracle_and_assign(d, nb, nlist, metric, xb, cent)
    g = _gpu(d, nlist, metric, cent, xb, assign, use_tc=use_tc, simt_max_avg_q=0.5 if use_tc else 3)
    g.nprobe = nprobe
    xq = clustered(nq, d, 4 * nlist, 12, normalize=(metric == IP))
    Dq, Iq = o.quantize(xq, nprobe)
    Do, Io = o.search_preassigned(xq, k, Iq)
    Dg, Ig = g.search_preassigned(xq, k, Iq)
    assert np.array_equal(Ig < 0, Io < 0), "padding differs: %d vs %d" % ((Ig < 0).sum(),
                                                                         (Io < 0).sum())
    assert_knn_parity(Do, Io, Dg, Ig, what="k=%d nprobe=%d" % (k, nprobe))


# ---- tensor-core path above 32 --------------------------------------------------
@pytest.mark.parametrize("metric", [L2, IP])
@pytest.mark.parametrize("nprobe", [48, 64, 128, 240])
def test_tc_quantizer_large_nprobe(cuda_device, metric, nprobe):
    import hnsw_ivf_b200 as b2
    from oracle import pyoracle as po
    d, nlist, nq = 96, 4096, 1500
    cent = clustered(nlist, d, 512, 3, normalize=(metric == IP))
    xq = clustered(nq, d, 512, 8, normalize=(metric == IP))
    g_tc, g_ex = b2.IndexIVFFlat(d, nlist, metric), b2.IndexIVFFlat(d, nlist, metric)
    g_ex.set_option("use_tc", 0)
    g_tc.set_centroids(cent)
    g_ex.set_centroids(cent)
    n0 = b2.tensor_core_launches()
    Dt, It = g_tc.quantize(xq, nprobe)
    n1 = b2.tensor_core_launches()
    Ds, Is = g_ex.quantize(xq, nprobe)
    n2 = b2.tensor_core_launches()
    assert n1 > n0 and n2 == n1, "tensor-core launches: %d / %d" % (n1 - n0, n2 - n1)
    assert np.array_equal(It, Is) and np.array_equal(Dt, Ds)
    Dr, Ir = po.port_knn(metric, xq, cent, nprobe)
    r = compare_knn(Dr, Ir, Dt, It, rtol=1e-4, atol=1e-4)
    assert r["unexplained"] == 0 and r["bad_dist"] == 0, r


@pytest.mark.parametrize("metric", [L2, IP])
@pytest.mark.parametrize("d,k,nprobe", [(64, 64, 8), (128, 100, 16), (96, 200, 8), (128, 40, 40)])
def test_tc_scan_large_k(cuda_device, metric, d, k, nprobe):
    import hnsw_ivf_b200 as b2
    nb, nlist, nq = 50000, 64, 1200
    xb = clustered(nb, d, 4 * nlist, 21, normalize=(metric == IP))
    cent = sample_centroids(xb, nlist)
    o, assign = _oracle_and_assign(d, nb, nlist, metric, xb, cent)
    g_tc = _gpu(d, nlist, metric, cent, xb, assign, simt_max_avg_q=0.5)
    g_ex = _gpu(d, nlist, metric, cent, xb, assign, use_tc=0)
    g_tc.nprobe = g_ex.nprobe = nprobe
    xq = clustered(nq, d, 4 * nlist, 22, normalize=(metric == IP))
    Dq, Iq = o.quantize(xq, nprobe)
    Do, Io = o.search_preassigned(xq, k, Iq)
    n0 = b2.tensor_core_launches()
    Dt, It = g_tc.search_preassigned(xq, k, Iq)
    n1 = b2.tensor_core_launches()
    Ds, Is = g_ex.search_preassigned(xq, k, Iq)
    assert n1 > n0 and b2.tensor_core_launches() == n1
    assert np.array_equal(It, Is) and np.array_equal(Dt, Ds)
    assert_knn_parity(Do, Io, Dt, It, what="tc k=%d" % k)
    De, Ie = g_tc.search(xq, k)  # coarse quantizer at nprobe 40 on the tensor cores too
    Df, If = g_ex.search(xq, k)
    assert np.array_equal(Ie, If) and np.array_equal(De, Df)


# ---- ADVICE (medium): near-ties under the 3xTF32 selection --------------------------------------
@pytest.mark.parametrize("k,nprobe", [(32, 32), (10, 8), (24, 24)])
def test_integer_sift_like_near_ties(cuda_device, k, nprobe):
    """SIFT-magnitude integer vectors (0..218, d = 128): exact L2 distances are integers ~1e5 with
    many rows 1..4 apart near the k-th boundary, while the |x|^2+|y|^2-2xy form of the tensor-core
    score carries ~0.5 of absolute error at |x||y| ~ 3e5. The rows that decide the result must
    still reach the exact re-rank: ids equal the exact kernel's, and the oracle's outside exact
    ties. Near-duplicates (a base row +-1 in one coordinate, 60 copies) put 60 rows within 1.0 of
    each other at the very top."""
    import hnsw_ivf_b200 as b2
    d, nb, nlist, nq = 128, 60000, 32, 600
    rng = np.random.default_rng(5)
    base = rng.gamma(1.2, 22.0, size=(nlist * 8, d)).clip(0, 218)
    which = rng.integers(0, base.shape[0], size=nb)
    xb = np.rint(base[which] + rng.normal(0, 9.0, size=(nb, d))).clip(0, 218).astype(np.float32)
    for j in range(60):  # near-duplicates of row 0
        xb[1000 + j] = xb[0]
        xb[1000 + j, j % d] += 1.0 if xb[0, j % d] < 200 else -1.0
    cent = sample_centroids(xb, nlist)
    o, assign = _oracle_and_assign(d, nb, nlist, L2, xb, cent)
    g_tc = _gpu(d, nlist, L2, cent, xb, assign, simt_max_avg_q=0.5)
    g_ex = _gpu(d, nlist, L2, cent, xb, assign, use_tc=0)
    g_tc.nprobe = g_ex.nprobe = nprobe
    xq = np.rint(base[rng.integers(0, base.shape[0], size=nq)] +
                 rng.normal(0, 9.0, size=(nq, d))).clip(0, 218).astype(np.float32)
    xq[:40] = xb[0]  # queries sitting on the near-duplicate cluster
    Dq, Iq = o.quantize(xq, nprobe)
    Do, Io = o.search_preassigned(xq, k, Iq)
    n0 = b2.tensor_core_launches()
    Dt, It = g_tc.search_preassigned(xq, k, Iq)
    assert b2.tensor_core_launches() > n0
    Ds, Is = g_ex.search_preassigned(xq, k, Iq)
    # integer distances are exact in fp32: the three tables agree to the bit, and ids may differ
    # only inside groups of EXACTLY equal distance (which member of a tie group that straddles the
    # k-th slot survives is the one thing the approximate selection may decide differently)
    assert np.array_equal(Dt, Ds), "distances differ from the exact kernel's"
    assert np.array_equal(Do, Dt), "distances differ from the oracle's"
    for Iother, name in ((Is, "exact kernel"), (Io, "oracle")):
        for q, j in zip(*np.nonzero(Iother != It)):
            same = Dt[q] == Dt[q, j]
            assert same.sum() > 1 or j == k - 1, "id differs from the %s at (%d,%d) outside a " \
                                                 "tie" % (name, q, j)
        # and never by more than a tie group: as multisets of (distance) the rows are identical,
        # every returned id really has the distance it is listed with
    chk = np.random.default_rng(1).integers(0, nq, size=50)
    for q in chk:
        for j in range(k):
            dd = float(((xb[It[q, j]] - xq[q]) ** 2).sum())
            assert dd == Dt[q, j], (q, j, dd, Dt[q, j])
    # the coarse quantizer at nprobe = 32 (KP = 64 on the tensor cores)
    Dc, Ic = g_tc.quantize(xq, nprobe)
    De, Ie = g_ex.quantize(xq, nprobe)
    assert np.array_equal(Ic, Ie) and np.array_equal(Dc, De)


# ---- reservations / in-place append --------------------------------------------------------------
@pytest.mark.parametrize("metric", [L2, IP])
def test_reserve_lists_in_place_append(cuda_device, metric):
    import hnsw_ivf_b200 as b2
    d, nb, nlist, k, nprobe = 48, 30000, 40, 10, 6
    xb = clustered(nb, d, 4 * nlist, 31, normalize=(metric == IP))
    cent = sample_centroids(xb, nlist)
    ids = np.arange(nb, dtype=np.int64) * 3 + 7
    whole = b2.IndexIVFFlat(d, nlist, metric)
    whole.set_centroids(cent)
    whole.add_with_ids(xb, ids)
    sizes = whole.list_sizes()
    assert sizes.sum() == nb and sizes[5] == whole.get_list_size(5)
    variants = []
    for reserve in (True, False):
        g = b2.IndexIVFFlat(d, nlist, metric)
        g.set_centroids(cent)
        if reserve:
            g.reserve_lists(sizes)
        for i0 in range(0, nb, 7001):  # chunks of odd size: lists fill across many calls
            g.add_with_ids(xb[i0:i0 + 7001], ids[i0:i0 + 7001])
        variants.append(g)
    xq = clustered(500, d, 4 * nlist, 32, normalize=(metric == IP))
    whole.nprobe = nprobe
    Dw, Iw = whole.search(xq, k)
    for g in variants:
        assert g.ntotal == nb and np.array_equal(g.list_sizes(), sizes)
        for l in (0, 7, nlist - 1):
            cw, iw = whole.get_list(l)
            cg, ig = g.get_list(l)
            assert np.array_equal(iw, ig) and np.array_equal(cw, cg)
        g.nprobe = nprobe
        Dg, Ig = g.search(xq, k)
        assert np.array_equal(Iw, Ig) and np.array_equal(Dw, Dg)
    # a reservation that is outgrown still works (the slab is re-laid out)
    g = b2.IndexIVFFlat(d, nlist, metric)
    g.set_centroids(cent)
    g.reserve_lists(np.maximum(sizes // 3, 1))
    g.add_with_ids(xb, ids)
    g.nprobe = nprobe
    Dg, Ig = g.search(xq, k)
    assert np.array_equal(Iw, Ig) and np.array_equal(Dw, Dg)


# ---- asynchronous all-device calls ---------------------------------------------------------------
def test_async_device_search_and_deferred_error(cuda_device):
    import torch
    import hnsw_ivf_b200 as b2
    d, nb, nlist, k, nprobe, nq = 64, 40000, 64, 10, 8, 3000
    xb = clustered(nb, d, 4 * nlist, 41)
    cent = sample_centroids(xb, nlist)
    g = b2.IndexIVFFlat(d, nlist, L2)
    g.set_centroids(cent)
    g.add(xb)
    g.nprobe = nprobe
    xq = clustered(nq, d, 4 * nlist, 42)
    Dh, Ih = g.search(xq, k)  # host buffers: synchronous
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        g.set_stream(st.cuda_stream)
        xd = torch.from_numpy(xq).cuda()
        outs = []
        g.reset_stats()
        for _ in range(20):  # more calls in flight than the handle has event sets
            outs.append(g.search(xd, k))
        st.synchronize()
        for Dd, Id in outs:
            assert np.array_equal(Id.cpu().numpy(), Ih) and np.array_equal(Dd.cpu().numpy(), Dh)
        calls, sums = g.stage_time_sums()
        assert calls == 20 and sums["total"] > 0 and sums["scan"] > 0
        assert sums["total"] >= sums["scan"] + sums["coarse"]
        t = g.last_timings()
        assert t["total"] > 0
        assert g.get_stats().nq == 20 * nq
        # an invalid key in a device-resident table is seen by the device only: reported late
        Iq = torch.full((nq, nprobe), 3, dtype=torch.int64, device="cuda")
        Iq[5, 2] = nlist + 4
        g.search_preassigned(xd, k, Iq)
        with pytest.raises(b2.B200Error) as ei:
            g.synchronize()
        assert "Invalid key" in str(ei.value)
        g.synchronize()  # reported once
        Dd, Id = g.search(xd, k)
        st.synchronize()
        assert np.array_equal(Id.cpu().numpy(), Ih)
    g.set_stream(None)


# ---- smaller advisor items -----------------------------------------------------------------------
def test_train_with_nlist_points_copies_them(cuda_device):
    """faiss/Clustering.cpp:346-368: nx == k -> the training vectors are the centroids, in order"""
    import hnsw_ivf_b200 as b2
    for metric in (L2, IP):
        d, nlist = 24, 50
        x = clustered(nlist, d, 16, 51)
        g = b2.IndexIVFFlat(d, nlist, metric)
        g.train(x)
        assert g.is_trained and np.array_equal(g.get_centroids(), x)


@pytest.mark.parametrize("metric", [L2, IP])
@pytest.mark.parametrize("nshard,k", [(8, 100), (3, 10), (16, 40)])
def test_merge_tie_order_is_by_shard_in_both_kernels(cuda_device, metric, nshard, k):
    """identical inputs through the rank merge (nshard * k <= 512) and the head walk (above) give
    the reference's order: equal distances lower shard first (faiss/utils/Heap.cpp:186-225)"""
    import hnsw_ivf_b200 as b2
    from oracle import pyoracle as po
    rng = np.random.default_rng(nshard + k)
    n = 130
    all_D = np.empty((nshard, n, k), np.float32)
    all_I = np.empty((nshard, n, k), np.int64)
    for s in range(nshard):
        v = rng.integers(0, 6, size=(n, k)).astype(np.float32)  # six values: ties everywhere
        v.sort(axis=1)
        all_D[s] = v if metric == L2 else -v
        all_I[s] = (rng.permutation(n * k).reshape(n, k) * nshard + s).astype(np.int64)
    Dg, Ig = b2.merge_knn_results(metric, all_D, all_I)
    Dp, Ip = po.port_merge_knn_results(metric, all_D, all_I)
    assert np.array_equal(Dg, Dp) and np.array_equal(Ig, Ip)


# ---- the sharded index over NCCL (needs >= 2 devices) ---------------------------------------------
def _nccl_rank(rank, world, port, q):
    try:
        sys.path.insert(0, ROOT)
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import torch
        import torch.distributed as dist
        import hnsw_ivf_b200 as b2
        from helpers import clustered, sample_centroids
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % port, rank=rank,
                                world_size=world, device_id=torch.device("cuda", rank))
        out = {}
        for metric in (L2, IP):
            d, nb, nlist, k, nq = 96, 120000, 256, 10, 2000
            xb = clustered(nb, d, 4 * nlist, 61, normalize=(metric == IP))
            cent = sample_centroids(xb, nlist)
            xq = clustered(nq, d, 4 * nlist, 62, normalize=(metric == IP))
            ids = np.arange(nb, dtype=np.int64) + 1000
            sh = b2.IndexShardsIVFFlat(d, nlist, metric, device=rank)
            sh.set_centroids(cent if rank == 0 else None, root=0)
            # each rank brings ITS slice of the vectors (uneven on purpose); the library routes
            # them to the owners of their lists
            cut = nb // 3
            lo, hi = (0, cut) if rank == 0 else (cut, nb)
            if world > 2:
                edges = np.linspace(0, nb, world + 1).astype(int)
                lo, hi = edges[rank], edges[rank + 1]
            sh.add_with_ids(xb[lo:hi], ids[lo:hi])
            assert sh.ntotal == nb
            single = b2.IndexIVFFlat(d, nlist, metric, device=rank)
            single.set_centroids(cent)
            single.add_with_ids(xb, ids)
            res = {}
            for nprobe in (1, 8, 40):
                sh.nprobe = single.nprobe = nprobe
                Ds, Is = single.search(xq, k)
                # (a) replicated host queries, (b) device queries broadcast from rank 0
                Dm, Im = sh.search(xq, k)
                xd = torch.from_numpy(xq).cuda() if rank == 0 else torch.zeros(
                    (nq, d), dtype=torch.float32, device="cuda")
                Db, Ib = sh.search(xd, k, broadcast_from=0)
                res[nprobe] = (bool(np.array_equal(Im, Is)), bool(np.array_equal(Dm, Ds)),
                               bool(np.array_equal(Ib.cpu().numpy(), Is)),
                               bool(np.array_equal(Db.cpu().numpy(), Ds)))
            calls, ph = sh.phase_time_sums()
            sizes = sh.local.list_sizes()
            out[metric] = (res, calls, ph["total"] > 0, int(sizes.sum()),
                           int(sizes[:sh.list_lo].sum() + sizes[sh.list_hi:].sum()))
            del sh, single
        dist.destroy_process_group()
        q.put((rank, out))
    except Exception:  # pragma: no cover
        import traceback
        q.put((rank, "ERR " + traceback.format_exc()))
        raise


def test_sharded_index_equals_single_nccl(cuda_device):
    """do_test_sharded_ivf (faiss/gpu/test/test_multi_gpu.py:54-96): the list-sharded index with a
    common quantizer returns the single index's table -- assert_array_equal on ids, and here on the
    distance bits too (the same kernels score the same rows)"""
    import torch
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs >= 2 CUDA devices")
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_nccl_rank, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in procs]
    for p in procs:
        p.join(timeout=120)
    total = 0
    for rank, out in res:
        assert not isinstance(out, str), out
        for metric, (per_nprobe, calls, timed, stored, foreign) in out.items():
            for nprobe, flags in per_nprobe.items():
                assert all(flags), (rank, metric, nprobe, flags)
            assert calls == 6 and timed and foreign == 0
            total += stored
    assert total == 2 * 120000  # every vector on exactly one rank, for both metrics
