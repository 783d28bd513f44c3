"""GPU tests of the ONE-MMA filter mode of the rows-major tensor-core scan (set_option("tc_split", 1)).

In that mode the tensor core computes tf32(x) . tf32(y) alone -- scores good to ~1.6e-3 |x||y| -- and
the k-select PROVES per query that the exactly re-scored candidates contain the exact top k
(kernels.h, SelectArgs::err_coef); a query it cannot prove is searched again on the exact warp
kernel. So the claim tested here is the strongest one: results are IDENTICAL, ids and distance
bits, to those of the SIMT kernels (B200IVF_NO_TC=1) -- on benign data, where the second pass never
runs, and on data built so that it must (hundreds of rows inside the error window of the k-th
distance)."""
import os

import numpy as np
import pytest

from helpers import clustered, sample_centroids

pytestmark = pytest.mark.gpu
L2, IP = 1, 0


def _handle(d, nlist, metric, tc):
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


def _pair(xb, cent, metric, assign=None, margin=6):
    d, nlist = xb.shape[1], cent.shape[0]
    hs = []
    for tc in (True, False):
        g = _handle(d, nlist, metric, tc)
        if tc:
            g.set_option("simt_max_avg_q", 0.5)  # tensor cores at every shape
            g.set_option("tc_mode", 2)           # rows-major kernel
            g.set_option("tc_split", 1)          # one MMA + checked re-rank
            g.set_option("one_margin", margin)
            g.set_option("one_min_items", 0)     # (small test shapes: few work items)
        g.set_centroids(cent)
        if assign is None:
            g.add_with_ids(xb, np.arange(xb.shape[0], dtype=np.int64))
        else:
            g.add_core(xb, None, assign)
        hs.append(g)
    return hs


@pytest.mark.parametrize("metric", [L2, IP])
@pytest.mark.parametrize("d,nb,nlist,nprobe,k,nq", [
    (128, 40000, 32, 8, 10, 2000),   # ~500 queries per list: many query tiles per list
    (96, 30000, 64, 16, 10, 1500),   # Deep1B-shaped rows
    (100, 20000, 32, 8, 28, 700),    # nf4 = 25 (the last MMA k-step is half padding), KP = 64
    (20, 20000, 16, 4, 1, 600),      # one TMA stage per tile, k = 1
    (64, 30000, 48, 12, 100, 900),   # k = 100: KP = 128
    (128, 3000, 64, 64, 10, 400),    # short lists (under one tile), nprobe = nlist
    (128, 60000, 512, 4, 10, 300),   # ~2 queries per list: the two-CTAs-per-SM shape
    (128, 50000, 64, 8, 200, 500),   # k = 200: KP = 256
])
def test_one_mma_equals_simt(cuda_device, metric, d, nb, nlist, nprobe, k, nq):
    import hnsw_ivf_b200 as b2
    xb = clustered(nb, d, 4 * nlist, 7, normalize=(metric == IP))
    cent = sample_centroids(xb, nlist)
    # (tight clusters far from the origin: the window of 1.6e-3 (|x|^2 + |y|^2) is wide against the
    # in-cluster distances, so a wide margin is what keeps the second pass rare here)
    g1, gs = _pair(xb, cent, metric, margin=16)
    xq = clustered(nq, d, 4 * nlist, 4321, normalize=(metric == IP))
    g1.nprobe = gs.nprobe = nprobe
    n0 = b2.tensor_core_launches()
    D1, I1 = g1.search(xq, k)
    n1 = b2.tensor_core_launches()
    Ds, Is = gs.search(xq, k)
    assert n1 > n0, "the tensor-core scan did not run"
    assert np.array_equal(I1, Is), "ids differ between the one-MMA filter and the SIMT path"
    assert np.array_equal(D1, Ds), "distances differ between the one-MMA filter and the SIMT path"
    # benign data: the proof goes through for (nearly) every query
    assert g1.exact_redo_count() <= nq // 50, g1.exact_redo_count()
    assert g1.get_stats().ndis == gs.get_stats().ndis


@pytest.mark.parametrize("metric", [L2, IP])
def test_one_mma_crowded_window_is_redone_exactly(cuda_device, metric):
    """Every query has hundreds of rows whose distances differ by ~1e-6 relative (far inside the
    1.6e-3 window): the kept sets cannot hold the window, the proof fails, the second pass runs --
    and the answer is still the exact one, ties and near-ties included."""
    rng = np.random.default_rng(5)
    d, nlist, k, nq = 64, 8, 10, 256
    base = rng.standard_normal((nlist, d)).astype(np.float32) * 4.0
    rows = []
    for l in range(nlist):
        # 600 tiny perturbations of one point per list, plus exact duplicates
        pert = base[l] + rng.standard_normal((600, d)).astype(np.float32) * 3e-5
        rows.append(pert)
        rows.append(np.repeat(base[l][None], 40, axis=0))
    xb = np.concatenate(rows).astype(np.float32)
    cent = base.copy()
    g1, gs = _pair(xb, cent, metric)
    xq = (base[rng.integers(0, nlist, nq)] +
          rng.standard_normal((nq, d)).astype(np.float32) * 0.5).astype(np.float32)
    g1.nprobe = gs.nprobe = 4
    D1, I1 = g1.search(xq, k)
    Ds, Is = gs.search(xq, k)
    assert g1.exact_redo_count() > 0, "the crowded window should have defeated the proof"
    assert np.array_equal(I1, Is)
    assert np.array_equal(D1, Ds)


def test_one_mma_selector_and_max_codes(cuda_device):
    import hnsw_ivf_b200 as b2
    d, nb, nlist, nprobe, k, nq = 64, 30000, 32, 8, 10, 1200
    xb = clustered(nb, d, 4 * nlist, 3)
    cent = sample_centroids(xb, nlist)
    g1, gs = _pair(xb, cent, L2)
    xq = clustered(nq, d, 4 * nlist, 99)
    for kw in (dict(sel=b2.IDSelectorRange(1000, 20000)), dict(max_codes=3000),
               dict(sel=b2.IDSelectorBatch(np.arange(0, nb, 3, dtype=np.int64)), max_codes=5000)):
        p = b2.SearchParametersIVF(nprobe=nprobe, **kw)
        D1, I1 = g1.search(xq, k, params=p)
        Ds, Is = gs.search(xq, k, params=p)
        assert np.array_equal(I1, Is), kw
        assert np.array_equal(D1, Ds), kw


def test_one_mma_device_buffers_and_preassigned(cuda_device):
    import torch
    d, nb, nlist, nprobe, k, nq = 96, 40000, 64, 8, 10, 3000
    xb = clustered(nb, d, 4 * nlist, 11)
    cent = sample_centroids(xb, nlist)
    g1, gs = _pair(xb, cent, L2)
    xq = clustered(nq, d, 4 * nlist, 12)
    g1.nprobe = gs.nprobe = nprobe
    xq_d = torch.from_numpy(xq).cuda()
    D1, I1 = g1.search(xq_d, k)
    torch.cuda.synchronize()
    Ds, Is = gs.search(xq, k)
    assert np.array_equal(I1.cpu().numpy(), Is)
    assert np.array_equal(D1.cpu().numpy(), Ds)
    _, Iq = gs.quantize(xq, nprobe)
    D2, I2 = g1.search_preassigned(xq, k, Iq)
    D3, I3 = gs.search_preassigned(xq, k, Iq)
    assert np.array_equal(I2, I3) and np.array_equal(D2, D3)


def test_one_mma_handle_adapts_to_crowded_data(cuda_device):
    """Tight Gaussian clusters far from the origin: with 6 spare keys the window does not fit for a
    good share of the queries of the first batch; the handle then keeps 16 spare keys (and would
    go on to the 3xTF32 split), so later batches are proven almost entirely -- and every batch
    returns the exact result."""
    d, nb, nlist, nprobe, k, nq = 128, 40000, 32, 8, 10, 2000
    xb = clustered(nb, d, 4 * nlist, 7)
    cent = sample_centroids(xb, nlist)
    g1, gs = _pair(xb, cent, L2, margin=6)
    g1.nprobe = gs.nprobe = nprobe
    redo = []
    for seed in (1, 2, 3):
        xq = clustered(nq, d, 4 * nlist, 1000 + seed)
        D1, I1 = g1.search(xq, k)
        Ds, Is = gs.search(xq, k)
        assert np.array_equal(I1, Is) and np.array_equal(D1, Ds)
        redo.append(g1.exact_redo_count())
    first, later = redo[0], redo[2] - redo[1]
    assert first > nq // 64, redo          # the first batch paid for the narrow margin
    assert later <= first // 3, redo       # the handle widened it


@pytest.mark.parametrize("metric", [L2, IP])
@pytest.mark.parametrize("d,nlist,nprobe,nq", [
    (128, 4096, 32, 3000),    # the rows-major kernel takes the centroid table above nprobe 24
    (96, 8192, 64, 2000),
    (64, 2048, 128, 1000),
    (256, 1024, 8, 1500),     # d > 128: rows-major at any nprobe
])
def test_one_mma_coarse_quantizer_equals_simt(cuda_device, metric, d, nlist, nprobe, nq):
    """K1 in the checked one-MMA form: probe ids and centroid distances identical to the SIMT
    quantizer's (the flagged queries, if any, are quantized again exactly)."""
    import hnsw_ivf_b200 as b2
    rng = np.random.default_rng(21)
    cent = rng.standard_normal((nlist, d)).astype(np.float32)
    if metric == IP:
        cent /= np.linalg.norm(cent, axis=1, keepdims=True)
    xq = (cent[rng.integers(0, nlist, nq)] + 0.3 * rng.standard_normal((nq, d))).astype(np.float32)
    g1 = _handle(d, nlist, metric, True)
    gs = _handle(d, nlist, metric, False)
    for g in (g1, gs):
        g.set_centroids(cent)
    g1.set_option("coarse_one_min_nlist", 0)  # (the default keeps this form for long tables)
    n0 = b2.tensor_core_launches()
    D1, I1 = g1.quantize(xq, nprobe)
    assert b2.tensor_core_launches() > n0
    Ds, Is = gs.quantize(xq, nprobe)
    assert np.array_equal(I1, Is)
    assert np.array_equal(D1, Ds)
