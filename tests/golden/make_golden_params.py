"""Generates tests/golden/params/*.npz by running the UNMODIFIED reference (oracle/_ref) on small
seeded inputs, for the callers' side of the path:

    python tests/golden/make_golden_params.py

  cases (JSON)   list of {sel, max_codes, store_pairs}: sel is a selector spec whose arrays are
                 named entries of the same file (tests/helpers.py::resolve_spec)
  D_i, I_i, ndis_i   index.search_preassigned(xq, k, Iq, ..., store_pairs, params) of case i
                     faiss/IndexIVF.cpp:399-723, faiss/IndexIVFFlat.cpp:155-179
  range_cases (JSON), rl_i / rD_i / rI_i   index.range_search_preassigned(xq, radius, Iq, ...)
                     faiss/IndexIVF.cpp:1243-1399, faiss/IndexIVFFlat.cpp:181-201
  xt, trained    index.train(xt) centroids: faiss/IndexIVF.cpp:66-127 -> faiss/Clustering.cpp:296-600
                 (niter 10, seed 1234, 256 points per centroid; a second set with 20 points per
                 centroid exercises the subsampling permutation)
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from helpers import clustered, resolve_spec, sample_centroids  # noqa: E402
from oracle import pyoracle as po  # noqa: E402


def main():
    po.ref_set_threads(1)
    for name, metric in (("params_l2", 1), ("params_ip", 0)):
        d, nb, nq, nlist, nprobe, k = 32, 2000, 60, 16, 5, 10
        normalize = metric == 0
        xb = clustered(nb, d, 4 * nlist, 31, normalize=normalize)
        xq = clustered(nq, d, 4 * nlist, 32, normalize=normalize)
        cent = sample_centroids(xb, nlist)
        rng = np.random.default_rng(33)
        ids = rng.permutation(4 * nb)[:nb].astype(np.int64)
        r = po.RefIndex(d, nlist, metric)
        r.set_centroids(cent)
        r.add_with_ids(xb, ids)
        pos = {int(v): i for i, v in enumerate(ids)}
        assign = np.full(nb, -1, np.int64)
        for l in range(nlist):
            assign[[pos[int(v)] for v in r.get_list(l)[1]]] = l
        Dq, Iq = r.quantize(xq, nprobe)
        out = dict(xb=xb, ids=ids, centroids=cent, assign=assign, xq=xq, Iq=Iq,
                   meta=np.array([metric, d, nb, nq, nlist, nprobe, k], np.int64))
        out["batch0"] = ids[::3].copy()
        out["batch1"] = ids[1::2].copy()
        out["bitmap0"] = rng.integers(0, 256, size=(4 * nb) // 8 - 40).astype(np.uint8)
        sels = [None, ["range", 500, 5000], ["batch", "batch0"], ["bitmap", "bitmap0"],
                ["not", ["range", 500, 5000]], ["and", ["range", 0, 6000], ["batch", "batch1"]],
                ["or", ["range", 0, 300], ["batch", "batch0"]],
                ["xor", ["bitmap", "bitmap0"], ["range", 1000, 7000]],
                ["range", 10**9, 10**9 + 5]]  # selects nothing
        cases = []
        for sp in sels:
            for mc in (0, 100, 333):
                cases.append(dict(sel=sp, max_codes=mc, store_pairs=False))
        for mc in (0, 100):
            cases.append(dict(sel=None, max_codes=mc, store_pairs=True))
        for i, c in enumerate(cases):
            po.ref_stats(reset=True)
            D, I = r.search_preassigned_ex(xq, k, Iq, max_codes=c["max_codes"],
                                           sel=resolve_spec(c["sel"], out),
                                           store_pairs=c["store_pairs"])
            out["D_%d" % i], out["I_%d" % i] = D, I
            out["ndis_%d" % i] = np.array([po.ref_stats()["ndis"]], np.int64)
        out["cases"] = np.array(json.dumps(cases))
        # radius: the median 5th-neighbour distance, so that most queries have a handful of hits
        D0 = out["D_0"]
        radius = float(np.median(D0[:, 4]))
        rcases = [dict(sel=None, store_pairs=False, radius=radius),
                  dict(sel=["range", 500, 5000], store_pairs=False, radius=radius),
                  dict(sel=["not", ["batch", "batch0"]], store_pairs=False, radius=radius),
                  dict(sel=None, store_pairs=True, radius=radius),
                  dict(sel=None, store_pairs=False,
                       radius=float(D0[:, 0].min()) if metric == 1 else float(D0[:, 0].max()))]
        for i, c in enumerate(rcases):
            l, D, I = r.range_search(xq, c["radius"], nprobe, keys=Iq,
                                     sel=resolve_spec(c["sel"], out), store_pairs=c["store_pairs"])
            out["rl_%d" % i], out["rD_%d" % i], out["rI_%d" % i] = l, D, I
        out["range_cases"] = np.array(json.dumps(rcases))
        # training
        xt = clustered(1500, d, 4 * nlist, 34, normalize=normalize)
        t = po.RefIndex(d, nlist, metric)
        t.train(xt)
        out["xt"] = xt
        out["trained"] = t.get_centroids()
        t2 = po.RefIndex(d, nlist, metric)
        t2.set_cp(niter=10, seed=1234, max_points_per_centroid=20)
        t2.train(xt)
        out["trained_sub20"] = t2.get_centroids()
        path = os.path.join(HERE, "params", name + ".npz")
        np.savez_compressed(path, **out)
        print("wrote", path, os.path.getsize(path) // 1024, "KB", len(cases), "cases")


if __name__ == "__main__":
    main()
