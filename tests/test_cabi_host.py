"""CPU tests of the host side: the C-ABI library loads and exports every symbol the header
declares, fails loudly without a GPU (no CPU fallback), and the list-sharded search logic
(IndexShardsIVF model, faiss/IndexShardsIVF.cpp:158-243) is right over a 2-process gloo group."""
import ctypes
import os
import re
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200ivf_c.h")


def _declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[A-Za-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import hnsw_ivf_b200 as b2
    from hnsw_ivf_b200 import binding
    lib = b2.load_library()
    declared = _declared_functions()
    assert len(declared) >= 25
    bound = {n for n, _, _ in binding.SYMBOLS}
    for name in declared:
        assert hasattr(lib, name), "libb200ivf.so lacks %s" % name
        assert name in bound, "binding.py does not bind %s" % name
    assert bound <= set(declared), "binding.py binds symbols the header does not declare"
    # raw dlopen too: the header is the contract, not the Python binding
    raw = ctypes.CDLL(b2.library_path())
    for name in declared:
        getattr(raw, name)


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import hnsw_ivf_b200 as b2
    with pytest.raises(b2.B200Error) as ei:
        b2.IndexIVFFlat(16, 4, b2.METRIC_L2)
    assert ei.value.code == -2 and "no CUDA device" in str(ei.value)
    assert b"no CUDA device" in b2.load_library().b200_get_last_error()


def test_enum_and_struct_layouts_match_reference_c_api():
    import hnsw_ivf_b200 as b2
    assert (b2.METRIC_INNER_PRODUCT, b2.METRIC_L2) == (0, 1)  # faiss/MetricType.h:23-24
    s = b2.IndexIVFStats()
    assert ctypes.sizeof(s) == 48  # c_api/IndexIVF_c.h:162-169: 4 x size_t + 2 x double
    assert [f[0] for f in s._fields_] == ["nq", "nlist", "ndis", "nheap_updates",
                                          "quantization_time", "search_time"]


def test_list_range_of_rank():
    import hnsw_ivf_b200 as b2
    for nlist, world in ((1024, 8), (10, 3), (7, 8), (262144, 8)):
        r = [b2.list_range_of_rank(nlist, i, world) for i in range(world)]
        assert r[0][0] == 0 and r[-1][1] == nlist
        assert all(r[i][1] == r[i + 1][0] for i in range(world - 1))


# ------------------------------------------------------------------------------------------------
# two ranks over gloo: each holds a list range, queries broadcast, tables gathered and merged
# ------------------------------------------------------------------------------------------------
class _CpuLocal:
    """stand-in for the device index with the same surface, backed by the oracle"""

    def __init__(self, d, nlist, metric):
        from oracle import pyoracle as po
        self.o = po.PortIndex(d, nlist, metric)
        self.nlist, self.nprobe, self.lo, self.hi = nlist, 1, 0, nlist

    def set_list_range(self, lo, hi):
        self.lo, self.hi = lo, hi

    def set_centroids(self, c):
        self.o.set_centroids(np.asarray(c))

    def add_with_ids(self, x, ids):
        x = np.asarray(x)
        _, a = self.o.quantize(x, 1)
        a = a[:, 0].copy()
        a[(a < self.lo) | (a >= self.hi)] = -1  # vectors of other ranks' lists are dropped
        self.o.add_core(x, ids, a)

    def search(self, x, k):
        import torch
        xn = x.numpy()
        _, Iq = self.o.quantize(xn, self.nprobe)
        return self.search_preassigned(x, k, torch.from_numpy(Iq))

    def quantize(self, x, nprobe):
        import torch
        Dq, Iq = self.o.quantize(np.ascontiguousarray(x.numpy()), nprobe)
        return torch.from_numpy(Dq), torch.from_numpy(Iq)

    def search_preassigned(self, x, k, Iq):
        import torch
        Iq = Iq.numpy()
        Iq = np.where((Iq >= self.lo) & (Iq < self.hi), Iq, -1)  # other ranks' lists: empty here
        D, I = self.o.search_preassigned(x.numpy(), k, np.ascontiguousarray(Iq))
        return torch.from_numpy(D), torch.from_numpy(I)


def _rank_main(rank, world, port, q):
    try:
        sys.path.insert(0, ROOT)
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import torch
        import torch.distributed as dist
        from helpers import clustered, sample_centroids
        from oracle import pyoracle as po
        import hnsw_ivf_b200 as b2
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        dist.init_process_group("gloo", rank=rank, world_size=world)
        out = {}
        for metric in (1, 0):
            d, nb, nq, nlist, nprobe, k = 16, 3000, 50, 12, 5, 10
            xb = clustered(nb, d, 48, 1)
            cent = sample_centroids(xb, nlist)
            xq = torch.from_numpy(clustered(nq, d, 48, 2))
            merge = lambda m, aD, aI: tuple(map(torch.from_numpy, po.port_merge_knn_results(
                m, aD.numpy(), aI.numpy())))
            sh = b2.ShardedSearchModel(d, nlist, metric, local=_CpuLocal(d, nlist, metric),
                                       merge=merge)
            sh.set_centroids(cent)
            sh.add_with_ids(xb, None)
            sh.nprobe = nprobe
            xin = xq.clone() if rank == 0 else torch.zeros_like(xq)
            D, I = sh.search(xin, k, broadcast_from=0)
            single = po.PortIndex(d, nlist, metric)
            single.set_centroids(cent)
            single.add_with_ids(xb)
            Ds, Is = single.search(xq.numpy(), k, nprobe)
            out[metric] = (bool(np.array_equal(I.numpy(), Is)), bool(np.array_equal(D.numpy(), Ds)),
                           sh.local.o.ntotal, (sh.list_lo, sh.list_hi))
        dist.destroy_process_group()
        q.put((rank, out))
    except Exception as e:  # pragma: no cover
        import traceback
        q.put((rank, "ERR " + traceback.format_exc()))
        raise e


def test_list_sharded_search_two_ranks_gloo():
    """tests/test_meta_index.py:128-176 / faiss/gpu/test/test_multi_gpu.py:54-96: shards + merge
    == the single index, over a real 2-process group"""
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, out in res:
        assert not isinstance(out, str), out
        for metric, (ids_eq, dis_eq, ntotal, rng) in out.items():
            assert ids_eq and dis_eq, (rank, metric)
            assert ntotal == 3000
