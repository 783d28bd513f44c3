"""CPU tests (no GPU) for the callers' side of the path: the C restatement
against golden vectors the UNMODIFIED reference produced (tests/golden/make_golden_params.py) --
SearchParametersIVF (IDSelector family, max_codes), store_pairs, range search, k-means training --
and, when oracle/_ref is present, against the compiled reference run live."""
import glob
import json
import os

import numpy as np
import pytest

from helpers import compare_knn, compare_range, resolve_spec
from oracle import pyoracle as po

PARAMS = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "params", "*.npz")))


def _load(path):
    g = dict(np.load(path))
    metric, d, nb, nq, nlist, nprobe, k = [int(v) for v in g["meta"]]
    o = po.PortIndex(d, nlist, metric)
    o.set_centroids(g["centroids"])
    o.add_core(g["xb"], g["ids"], g["assign"])
    return g, o, (metric, d, nb, nq, nlist, nprobe, k)


@pytest.mark.parametrize("path", PARAMS, ids=[os.path.basename(p) for p in PARAMS])
def test_port_search_params_match_reference_golden(path):
    g, o, (metric, d, nb, nq, nlist, nprobe, k) = _load(path)
    cases = json.loads(str(g["cases"]))
    assert len(cases) >= 20
    for i, c in enumerate(cases):
        o.stats.ndis = 0
        D, I = o.search_preassigned_ex(g["xq"], k, g["Iq"], max_codes=c["max_codes"],
                                       sel=resolve_spec(c["sel"], g), store_pairs=c["store_pairs"])
        r = compare_knn(g["D_%d" % i], g["I_%d" % i], D, I)
        assert r["unexplained"] == 0 and r["bad_dist"] == 0 and r["exact_id_rate"] > 0.999, (c, r)
        assert o.stats.ndis == int(g["ndis_%d" % i][0]), c  # max_codes accounting (IndexIVF.cpp:624)


@pytest.mark.parametrize("path", PARAMS, ids=[os.path.basename(p) for p in PARAMS])
def test_port_range_search_matches_reference_golden(path):
    g, o, (metric, d, nb, nq, nlist, nprobe, k) = _load(path)
    for i, c in enumerate(json.loads(str(g["range_cases"]))):
        l, D, I = o.range_search_preassigned(g["xq"], c["radius"], g["Iq"],
                                             sel=resolve_spec(c["sel"], g),
                                             store_pairs=c["store_pairs"])
        nb_ = compare_range(g["rl_%d" % i], g["rD_%d" % i], g["rI_%d" % i], l, D, I, c["radius"],
                            metric)
        assert nb_ <= 2, (c, nb_)
        if nb_ == 0:  # same sets: then also the reference's order (probe order, then list order)
            assert np.array_equal(l, g["rl_%d" % i]) and np.array_equal(I, g["rI_%d" % i]), c


@pytest.mark.parametrize("path", PARAMS, ids=[os.path.basename(p) for p in PARAMS])
def test_port_training_matches_reference_golden(path):
    """same subsample, same initial picks (MT19937 restated), same summation order: the centroids
    agree to rounding unless an assignment flips on a near-tie (none at this size)"""
    g, o, (metric, d, nb, nq, nlist, nprobe, k) = _load(path)
    for key, mpc in (("trained", 256), ("trained_sub20", 20)):
        t = po.PortIndex(d, nlist, metric)
        t.train(g["xt"], niter=10, seed=1234, max_points_per_centroid=mpc)
        np.testing.assert_allclose(t.get_centroids(), g[key], rtol=0, atol=2e-6, err_msg=key)


def test_selector_predicates():
    """is_member of every selector kind == the pure-Python statement of IDSelector.cpp:17-123"""
    rng = np.random.default_rng(0)
    batch = rng.integers(0, 500, size=60)
    bitmap = rng.integers(0, 256, size=40).astype(np.uint8)
    specs = [("range", 10, 200), ("batch", batch), ("bitmap", bitmap),
             ("not", ("batch", batch)), ("and", ("range", 0, 300), ("bitmap", bitmap)),
             ("or", ("range", 400, 450), ("batch", batch)),
             ("xor", ("bitmap", bitmap), ("range", 100, 400))]
    L = po.port_lib()
    probe = list(range(-3, 520)) + [2**40, -2**40]
    for sp in specs:
        t = po.port_sel(sp)
        for i in probe:
            assert bool(L.orc_sel_is_member(t.h, i)) == po.sel_is_member_py(sp, i), (sp[0], i)
        if po.ref_available():
            rt = po.ref_sel(sp)
            R = po.ref_lib()
            for i in probe:
                assert bool(R.ref_sel_is_member(rt.h, i)) == po.sel_is_member_py(sp, i), (sp[0], i)


@pytest.mark.skipif(not po.ref_available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("metric", [1, 0])
def test_port_matches_live_reference_params(metric):
    """fresh seeds, the reference run live: end-to-end search with params (coarse quantizer
    included), max_codes on the index field, the fork's search_stats results"""
    from helpers import clustered, sample_centroids
    d, nb, nq, nlist, nprobe, k = 24, 3000, 50, 24, 6, 7
    xb = clustered(nb, d, 4 * nlist, 41, normalize=metric == 0)
    xq = clustered(nq, d, 4 * nlist, 42, normalize=metric == 0)
    cent = sample_centroids(xb, nlist)
    r = po.RefIndex(d, nlist, metric)
    r.set_centroids(cent)
    r.add_with_ids(xb)
    o = po.PortIndex(d, nlist, metric)
    o.set_centroids(cent)
    o.add_with_ids(xb)
    Dq, Iq = r.quantize(xq, nprobe)
    for sp, mc in ((("range", 100, 2000), 0), (None, 250), (("not", ("range", 0, 1500)), 400)):
        Dr, Ir = r.search_params(xq, k, nprobe, max_codes=mc, sel=sp)
        Dp, Ip = o.search_preassigned_ex(xq, k, Iq, max_codes=mc, sel=sp)
        c = compare_knn(Dr, Ir, Dp, Ip)
        assert c["unexplained"] == 0 and c["bad_dist"] == 0, (sp, mc, c)
    Ds, Is, st = r.search_stats(xq, k, nprobe)
    Dp, Ip = o.search_preassigned(xq, k, Iq)
    c = compare_knn(Ds, Is, Dp, Ip)
    assert c["unexplained"] == 0 and c["bad_dist"] == 0, c
    assert st.shape == (nq, 3) and np.all(st[:, 0] >= st[:, 2])
