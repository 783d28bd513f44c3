"""GPU tests of the tensor-core scan (csrc/scan_tc.cu, tcgen05 3xTF32) specifically.

The tensor core only SELECTS candidates; every distance that leaves the library is re-scored in
plain fp32 by the k-select. So the strongest statement available is tested: with the tensor-core
path on and off (B200IVF_NO_TC=1 -> SIMT kernels only) the results are IDENTICAL, bit for bit,
and both agree with the CPU oracle under the tie-aware comparison of the other GPU tests."""
import os

import numpy as np
import pytest

from helpers import assert_knn_parity, clustered, compare_knn, sample_centroids

pytestmark = pytest.mark.gpu
L2, IP = 1, 0


def _handle(d, nlist, metric, tc):
    """device index with the tensor-core scan enabled / disabled (read at construction)"""
    import hnsw_ivf_b200 as b2
    old = os.environ.get("B200IVF_NO_TC")
    os.environ["B200IVF_NO_TC"] = "0" if tc else "1"
    try:
        return b2.IndexIVFFlat(d, nlist, metric)
    finally:
        if old is None:
            del os.environ["B200IVF_NO_TC"]
        else:
            os.environ["B200IVF_NO_TC"] = old


def _pair(d, nb, nlist, metric, seed=1):
    from oracle import pyoracle as po
    xb = clustered(nb, d, 4 * nlist, seed, normalize=(metric == IP))
    cent = sample_centroids(xb, nlist)
    o = po.PortIndex(d, nlist, metric)
    o.set_centroids(cent)
    o.add_with_ids(xb)
    assign = np.full(nb, -1, np.int64)
    for l in range(nlist):
        assign[o.get_list(l)[1]] = l
    hs = []
    for tc in (True, False):
        g = _handle(d, nlist, metric, tc)
        if tc:  # these tests are about the tensor-core kernels: keep them on at every shape
            g.set_option("simt_max_avg_q", 0.5)
        g.set_centroids(cent)
        g.add_core(xb, None, assign)
        hs.append(g)
    return xb, cent, o, hs[0], hs[1]


@pytest.mark.parametrize("metric", [L2, IP])
@pytest.mark.parametrize("d,nb,nlist,nprobe,k,nq", [
    (128, 40000, 32, 8, 10, 2000),   # ~500 queries per list: 4 query tiles per list, long lists
    (96, 30000, 64, 16, 10, 1500),   # Deep1B-shaped rows
    (100, 20000, 32, 8, 28, 700),    # nf4 = 25 (odd: the last MMA k-step is half padding), KP = 32
    (20, 20000, 16, 4, 1, 600),      # one TMA stage per tile, k = 1
    (64, 30000, 48, 12, 32, 900),    # largest k the tensor-core path takes
    (128, 3000, 64, 64, 10, 400),    # short lists (under one 128-row tile), nprobe = nlist
    (128, 60000, 512, 4, 10, 300),   # ~2 queries per list: rows-major kernel, N = 16 tiles
    (960, 12000, 32, 8, 10, 500),    # GIST-shaped rows (C3): rows-major kernel, 16 queries per item
    (200, 20000, 32, 8, 10, 700),    # d > 128 with an odd number of float4 columns (50)
    (260, 9000, 16, 4, 20, 300),     # d > 128, KP = 32
])
def test_tc_equals_simt_and_oracle(cuda_device, metric, d, nb, nlist, nprobe, k, nq):
    import hnsw_ivf_b200 as b2
    xb, cent, o, g_tc, g_simt = _pair(d, nb, nlist, metric)
    xq = clustered(nq, d, 4 * nlist, 4321, normalize=(metric == IP))
    Dq, Iq = o.quantize(xq, nprobe)
    Do, Io = o.search_preassigned(xq, k, Iq)
    g_tc.nprobe = g_simt.nprobe = nprobe
    n0 = b2.tensor_core_launches()
    Dt, It = g_tc.search_preassigned(xq, k, Iq)
    n1 = b2.tensor_core_launches()
    Ds, Is = g_simt.search_preassigned(xq, k, Iq)
    n2 = b2.tensor_core_launches()
    assert n1 > n0 and n2 == n1, "the tensor-core scan ran %d / %d times" % (n1 - n0, n2 - n1)
    assert np.array_equal(It, Is), "ids differ between the tensor-core and the SIMT path"
    assert np.array_equal(Dt, Ds), "distances differ between the tensor-core and the SIMT path"
    assert_knn_parity(Do, Io, Dt, It, what="tc d=%d k=%d" % (d, k))
    # end to end (coarse quantizer on the tensor cores too): same probes => same results
    De, Ie = g_tc.search(xq, k)
    Df, If = g_simt.search(xq, k)
    assert np.array_equal(Ie, If) and np.array_equal(De, Df)


@pytest.mark.parametrize("metric", [L2, IP])
@pytest.mark.parametrize("nprobe", [1, 8, 32])
def test_tc_quantizer(cuda_device, metric, nprobe):
    from oracle import pyoracle as po
    d, nlist, nq = 96, 2048, 3000
    cent = clustered(nlist, d, 256, 3, normalize=(metric == IP))
    xq = clustered(nq, d, 256, 8, normalize=(metric == IP))
    g_tc, g_simt = _handle(d, nlist, metric, True), _handle(d, nlist, metric, False)
    g_tc.set_centroids(cent)
    g_simt.set_centroids(cent)
    Dt, It = g_tc.quantize(xq, nprobe)
    Ds, Is = g_simt.quantize(xq, nprobe)
    assert np.array_equal(It, Is) and np.array_equal(Dt, Ds)
    Dr, Ir = po.port_knn(metric, xq, cent, nprobe)
    r = compare_knn(Dr, Ir, Dt, It, rtol=1e-4, atol=1e-4)
    assert r["unexplained"] == 0 and r["bad_dist"] == 0, r


def test_tc_cancellation_and_ties(cuda_device):
    """Queries that ARE database vectors (L2 distance exactly 0: the |x|^2+|y|^2-2xy form of the
    tensor-core score cancels completely) and 40 exact duplicates of one vector (a tie group wider
    than k and than the kept margin)."""
    d, nb, nlist, nprobe, k = 128, 20000, 16, 4, 10
    xb, cent, o, g_tc, g_simt = _pair(d, nb, nlist, L2, seed=9)
    xq = np.ascontiguousarray(xb[:500])
    Dq, Iq = o.quantize(xq, nprobe)
    Do, Io = o.search_preassigned(xq, k, Iq)
    g_tc.nprobe = nprobe
    Dt, It = g_tc.search_preassigned(xq, k, Iq)
    assert np.array_equal(It[:, 0], np.arange(500)) and np.all(Dt[:, 0] == 0.0)
    assert_knn_parity(Do, Io, Dt, It, what="self queries")
    # duplicates
    import hnsw_ivf_b200 as b2
    from oracle import pyoracle as po
    xb2 = xb.copy()
    xb2[100:140] = xb2[100]
    o2 = po.PortIndex(d, nlist, L2)
    o2.set_centroids(cent)
    o2.add_with_ids(xb2)
    assign = np.full(nb, -1, np.int64)
    for l in range(nlist):
        assign[o2.get_list(l)[1]] = l
    g2 = b2.IndexIVFFlat(d, nlist, L2)
    g2.set_centroids(cent)
    g2.add_core(xb2, None, assign)
    g2.nprobe = nprobe
    xq2 = np.ascontiguousarray(np.repeat(xb2[100:101], 64, axis=0) +
                               1e-3 * clustered(64, d, 4, 5))
    Dq2, Iq2 = o2.quantize(xq2, nprobe)
    Do2, Io2 = o2.search_preassigned(xq2, k, Iq2)
    Dg2, Ig2 = g2.search_preassigned(xq2, k, Iq2)
    # all 10 results are members of the tie group, distances equal to the oracle's
    assert np.all((Ig2 >= 100) & (Ig2 < 140))
    r = compare_knn(Do2, Io2, Dg2, Ig2)
    assert r["bad_dist"] == 0 and r["unexplained"] == 0, r
    # cmp2 order inside the tie group: ascending id for L2 (faiss/utils/ordered_key_value.h:51-53)
    assert np.array_equal(Ig2, Io2)
