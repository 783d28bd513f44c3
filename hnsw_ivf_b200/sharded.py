"""List-sharded IVF-Flat over several B200s, one process per GPU.

``IndexShardsIVFFlat`` binds the C-ABI's ``b200_IndexShardsIVFFlat_*`` (csrc/shards.cu): the C++
object that owns this rank's shard and its NCCL communicator and runs a sharded search as ONE
stream of device work -- coarse quantizer split by query, int32 probe-table all-gather, local
plan + scan + k-select, one grouped all-gather of the per-shard tables, device merge. Behavioural
model: IndexShardsIVF (faiss/IndexShardsIVF.cpp:102-243) with shards that hold list RANGES
(shard_type 4 / SUBSET_TYPE_INVLIST: faiss/gpu/GpuCloner.cpp:303-314,
faiss/invlists/InvertedLists.cpp:168-176). Each (query, probe) pair lands on exactly one rank, so
total scan bytes do not grow with the number of GPUs. torch.distributed is used for ONE thing
here: handing rank 0's NCCL unique id to the other ranks.

``ShardedSearchModel`` is the same algorithm written over torch.distributed collectives with an
injectable local index: the host-logic model the CPU test suite runs over gloo with the oracle as
the local index (tests/test_cabi_host.py). It is not on the product path.
"""
import ctypes as C

import numpy as np

from .binding import (IndexIVFFlat, _empty_like_input, _is_torch, _ptr, _vp, load_library,
                      merge_knn_results, B200Error)

PHASES = ("broadcast", "coarse", "probe_gather", "local_search", "result_gather", "merge", "total")


def list_range_of_rank(nlist, rank, world):
    """Lists [lo, hi) owned by `rank`: same arithmetic as faiss/gpu/GpuCloner.cpp:304-305."""
    return (rank * nlist) // world, ((rank + 1) * nlist) // world


def nccl_unique_id():
    """128 bytes of ncclGetUniqueId (b200_nccl_get_unique_id)"""
    lib = load_library()
    buf = C.create_string_buffer(128)
    rc = lib.b200_nccl_get_unique_id(buf)
    if rc != 0:
        raise B200Error(rc, lib.b200_get_last_error().decode())
    return buf.raw


class IndexShardsIVFFlat:
    """This rank's end of a list-sharded IndexIVFFlat (every method is collective: all ranks call
    it with the same arguments, except where noted).

    ``unique_id``: the 128 NCCL id bytes shared by all ranks; by default rank 0 of the
    torch.distributed ``group`` makes one and broadcasts it.
    """

    def __init__(self, d, nlist, metric, device=0, group=None, rank=None, world=None,
                 unique_id=None):
        self._lib = load_library()
        self._h = None
        if rank is None or world is None:
            import torch.distributed as dist
            rank, world = dist.get_rank(group), dist.get_world_size(group)
        if unique_id is None and world > 1:
            import torch.distributed as dist
            box = [nccl_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group else 0,
                                       group=group)
            unique_id = box[0]
        h = _vp()
        idbuf = C.create_string_buffer(unique_id, 128) if unique_id is not None else None
        self._check(self._lib.b200_IndexShardsIVFFlat_new(C.byref(h), d, nlist, metric, device,
                                                          rank, world, idbuf))
        self._h = h
        self.d, self.nlist, self.metric, self.device = d, nlist, metric, device
        self.rank, self.world = rank, world
        self.list_lo, self.list_hi = list_range_of_rank(nlist, rank, world)
        self.local = IndexIVFFlat(d, nlist, metric, device,
                                  _handle=_vp(self._lib.b200_IndexShardsIVFFlat_local(h)),
                                  _borrowed=True)

    def __del__(self):
        if getattr(self, "_h", None):
            self._lib.b200_IndexShardsIVFFlat_free(self._h)
            self._h = None

    def _check(self, rc):
        if rc != 0:
            raise B200Error(rc, self._lib.b200_get_last_error().decode())

    @property
    def nprobe(self):
        return self.local.nprobe

    @nprobe.setter
    def nprobe(self, v):
        self.local.nprobe = v

    @property
    def ntotal(self):
        return self._lib.b200_IndexShardsIVFFlat_ntotal(self._h)

    def set_stream(self, cuda_stream_ptr):
        self.local.set_stream(cuda_stream_ptr)

    def set_centroids(self, c, root=-1):
        """root < 0: every rank passes the same table; root >= 0: rank root's is broadcast"""
        keep, p = _ptr(c, np.float32, (self.nlist, self.d), "centroids")
        self.local._order_with_torch(keep)
        self._check(self._lib.b200_IndexShardsIVFFlat_set_centroids(self._h, p, root))

    def train(self, x, root=0):
        kx, px = _ptr(x, np.float32, None, "x")
        self.local._order_with_torch(kx)
        n = kx.shape[0] if kx is not None else 0
        self._check(self._lib.b200_IndexShardsIVFFlat_train(self._h, n, px, root))

    def add_with_ids(self, x, ids):
        """NOT replicated: each rank passes ITS OWN vectors (possibly none); the library sends
        every vector to the rank that owns its list."""
        self.add_core(x, ids, None)

    def add_core(self, x, ids, assign):
        kx, px = _ptr(x, np.float32, None, "x")
        n = kx.shape[0] if kx is not None else 0
        ki, pi = _ptr(ids, np.int64, (n,), "ids")
        ka, pa = _ptr(assign, np.int64, (n,), "assign")
        self.local._order_with_torch(kx, ki, ka)
        if assign is None:
            self._check(self._lib.b200_IndexShardsIVFFlat_add_with_ids(self._h, n, px, pi))
        else:
            self._check(self._lib.b200_IndexShardsIVFFlat_add_core(self._h, n, px, pi, pa))

    def reserve_lists(self, sizes):
        self.local.reserve_lists(sizes)

    def search(self, x, k, D=None, I=None, broadcast_from=None):
        """x: [n][d], the same on every rank -- or, with ``broadcast_from=r``, rank r's queries
        (the other ranks pass a buffer of the same shape to receive them). Returns the merged
        (D, I) on every rank, living where x lives."""
        kx, px = _ptr(x, np.float32, None, "x")
        n = kx.shape[0]
        if kx.ndim != 2 or kx.shape[1] != self.d:
            raise ValueError("x must be [n][%d]" % self.d)
        if D is None:
            D = _empty_like_input(kx, (n, k), np.float32)
        if I is None:
            I = _empty_like_input(kx, (n, k), np.int64)
        kd, pd = _ptr(D, np.float32, (n, k), "D")
        ki, pi = _ptr(I, np.int64, (n, k), "I")
        self.local._order_with_torch(kx, kd, ki)
        root = -1 if broadcast_from is None else int(broadcast_from)
        self._check(self._lib.b200_IndexShardsIVFFlat_search(self._h, n, px, k, pd, pi, root))
        if self.local._own_stream and _is_torch(kx) and kx.is_cuda:
            self.synchronize()
        return kd, ki

    def synchronize(self):
        self._check(self._lib.b200_IndexShardsIVFFlat_synchronize(self._h))

    def reset_phase_times(self):
        self._lib.b200_IndexShardsIVFFlat_reset_phase_times(self._h)

    def phase_time_sums(self):
        """(calls, {phase: device ms summed over the searches since reset_phase_times})"""
        ms = (C.c_double * 7)()
        calls = self._lib.b200_IndexShardsIVFFlat_phase_time_sums(self._h, ms)
        return int(calls), dict(zip(PHASES, list(ms)))


# ------------------------------------------------------------------------------------------------
# host-logic model (CPU tests over gloo; not the product path)
# ------------------------------------------------------------------------------------------------
def gather_results(D_loc, I_loc, group=None):
    """All-gather the per-rank (n,k) tables into [world][n][k] tensors."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    n, k = D_loc.shape
    all_D = torch.empty((world * n, k), dtype=D_loc.dtype, device=D_loc.device)
    all_I = torch.empty((world * n, k), dtype=I_loc.dtype, device=I_loc.device)
    dist.all_gather_into_tensor(all_D, D_loc.contiguous(), group=group)
    dist.all_gather_into_tensor(all_I, I_loc.contiguous(), group=group)
    return all_D.view(world, n, k), all_I.view(world, n, k)


class ShardedSearchModel:
    """The algorithm of csrc/shards.cu over torch.distributed collectives, with the local index
    and the merge injected (the CPU suite uses the oracle for both, over a 2-process gloo group):
    same list ranges, same query slices of the common quantizer (padded rows = -1), same probe
    table on every rank, every pair scanned by the one rank that owns its list, all-gather, merge.
    Here every rank is SHOWN all the vectors and keeps those of its lists."""

    def __init__(self, d, nlist, metric, local, merge, group=None):
        import torch.distributed as dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.d, self.nlist, self.metric = d, nlist, metric
        self.list_lo, self.list_hi = list_range_of_rank(nlist, self.rank, self.world)
        local.set_list_range(self.list_lo, self.list_hi)
        self.local = local
        self._merge = merge

    @property
    def nprobe(self):
        return self.local.nprobe

    @nprobe.setter
    def nprobe(self, v):
        self.local.nprobe = v

    def set_centroids(self, c):
        self.local.set_centroids(c)

    def add_with_ids(self, x, ids):
        self.local.add_with_ids(x, ids)

    def search(self, x, k, broadcast_from=None):
        import torch
        import torch.distributed as dist
        if broadcast_from is not None:
            dist.broadcast(x, src=broadcast_from, group=self.group)
        Iq = self._quantize_sharded(x)
        D_loc, I_loc = self.local.search_preassigned(x, k, Iq)
        all_D, all_I = gather_results(torch.as_tensor(D_loc), torch.as_tensor(I_loc), self.group)
        return self._merge(self.metric, all_D, all_I)

    def _quantize_sharded(self, x):
        import torch
        import torch.distributed as dist
        n = x.shape[0]
        nprobe = min(self.local.nprobe, self.nlist)
        chunk = (n + self.world - 1) // self.world
        lo, hi = min(self.rank * chunk, n), min((self.rank + 1) * chunk, n)
        allq = torch.full((self.world * chunk, nprobe), -1, dtype=torch.int64, device=x.device)
        if hi > lo:
            _, Iq = self.local.quantize(x[lo:hi].contiguous(), nprobe)
            allq[self.rank * chunk:self.rank * chunk + hi - lo] = torch.as_tensor(Iq).to(x.device)
        mine = allq[self.rank * chunk:(self.rank + 1) * chunk].clone()
        dist.all_gather_into_tensor(allq, mine, group=self.group)
        return allq[:n]


__all__ = ["IndexShardsIVFFlat", "ShardedSearchModel", "list_range_of_rank", "nccl_unique_id",
           "merge_knn_results", "PHASES"]
