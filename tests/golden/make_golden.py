"""Generates tests/golden/*.npz by running the UNMODIFIED reference (oracle/_ref, compiled from
/root/reference by oracle/Makefile) on small seeded inputs. Run in the build container:

    python tests/golden/make_golden.py

Each file holds the inputs (so that the fixtures do not depend on a numpy generator version) and
the reference's outputs for one case of the path:
  centroids, xb, ids, xq                         inputs
  assign        list number the reference's add_with_ids gave every vector
  Dq, Iq        quantizer->search(nq, xq, nprobe)          faiss/IndexFlat.cpp:38
  D, I          index.search(nq, xq, k) with nprobe        faiss/IndexIVF.cpp:303
  Dp, Ip        index.search_preassigned(..., Iq, Dq)      faiss/IndexIVF.cpp:399
  ndis          indexIVF_stats.ndis of the search          faiss/IndexIVF.cpp:721
<name>.ivfflat (lowlevel_l2, lowlevel_ip, sparse_l2): the index file faiss::write_index wrote for
the case (faiss/impl/index_write.cpp:635-638) -- golden bytes for the file reader / writer.
The shapes follow the reference's own tests for this path (tests/test_lowlevel_ivf.cpp:43-75:
d=32, nb=1000, nq=200, nlist=32, nprobe=4, k=10) plus k>=100, d%4!=0, nprobe=nlist and ties.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from helpers import clustered, sample_centroids  # noqa: E402
from oracle import pyoracle as po  # noqa: E402

CASES = [
    # name, metric, d, nb, nq, nlist, nprobe, k, duplicate_every
    ("lowlevel_l2", 1, 32, 1000, 200, 32, 4, 10, 0),
    ("lowlevel_ip", 0, 32, 1000, 200, 32, 4, 10, 0),
    ("odd_d_l2", 1, 30, 3000, 64, 16, 3, 5, 0),
    ("k100_ip", 0, 64, 4000, 32, 20, 5, 100, 0),
    ("exhaustive_l2", 1, 16, 2000, 40, 8, 8, 10, 0),
    ("ties_l2", 1, 8, 600, 30, 4, 2, 12, 3),
    ("ties_ip", 0, 8, 600, 30, 4, 2, 12, 3),
    ("sift_shape_l2", 1, 128, 4000, 100, 32, 8, 10, 0),
]


INDEX_FILES = ("lowlevel_l2", "lowlevel_ip")


def main():
    po.ref_set_threads(1)
    for name, metric, d, nb, nq, nlist, nprobe, k, dup in CASES:
        normalize = metric == 0
        xb = clustered(nb, d, 4 * nlist, 11, normalize=normalize)
        if dup:  # exact duplicates: equal distances, exercises the cmp2 tie order
            xb[dup::dup] = xb[0:nb - dup:dup][: xb[dup::dup].shape[0]]
        xq = clustered(nq, d, 4 * nlist, 22, normalize=normalize)
        if dup:
            xq[:5] = xb[:5]
        cent = sample_centroids(xb, nlist)
        ids = (np.arange(nb, dtype=np.int64) * 7 + 3)
        r = po.RefIndex(d, nlist, metric)
        r.set_centroids(cent)
        r.add_with_ids(xb, ids)
        assign = np.full(nb, -1, np.int64)
        pos = {int(v): i for i, v in enumerate(ids)}
        for l in range(nlist):
            _, li = r.get_list(l)
            for v in li:
                assign[pos[int(v)]] = l
        Dq, Iq = r.quantize(xq, nprobe)
        po.ref_stats(reset=True)
        D, I = r.search(xq, k, nprobe)
        ndis = po.ref_stats(reset=True)["ndis"]
        Dp, Ip = r.search_preassigned(xq, k, Iq, Dq)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), metric=metric, nlist=nlist,
                            nprobe=nprobe, k=k, centroids=cent, xb=xb, ids=ids, xq=xq,
                            assign=assign, Dq=Dq, Iq=Iq, D=D, I=I, Dp=Dp, Ip=Ip, ndis=ndis)
        print(name, "ndis", ndis, "I[0]", I[0][:5])
        if name in INDEX_FILES:
            # the reference's own index file of this case (faiss::write_index), "full" size table
            r.set_nprobe(nprobe)
            r.write_index(os.path.join(HERE, name + ".ivfflat"))
    # an index with most lists empty: write_index stores the "sprs" size table
    # (faiss/impl/index_write.cpp:268-279)
    d, nlist, nb = 16, 40, 300
    xb = clustered(nb, d, 8, 5)
    cent = sample_centroids(xb, nlist)
    assign = np.array([3, 17, 38], np.int64)[np.arange(nb) % 3]
    ids = np.arange(nb, dtype=np.int64) + 1000
    r = po.RefIndex(d, nlist, 1)
    r.set_centroids(cent)
    r.add_core(xb, ids, assign)
    r.set_nprobe(5)
    r.write_index(os.path.join(HERE, "sparse_l2.ivfflat"))
    with open(os.path.join(HERE, "sparse_l2.inputs"), "wb") as f:  # (an .npz by content)
        np.savez_compressed(f, centroids=cent, xb=xb, ids=ids, assign=assign)


if __name__ == "__main__":
    main()
