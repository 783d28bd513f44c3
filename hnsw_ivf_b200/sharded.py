"""List-sharded IVF-Flat over several B200s, one process per GPU (torch.distributed / NCCL).

Behavioural model: IndexShardsIVF::search (faiss/IndexShardsIVF.cpp:158-243) -- a common coarse
quantizer, one search_preassigned per shard, merge_knn_results -- with the reference's host threads
and CPU merge replaced by: replicated device quantizer, contiguous list ranges per rank
(shard_type 4 / SUBSET_TYPE_INVLIST: faiss/gpu/GpuCloner.cpp:303-314,
faiss/invlists/InvertedLists.cpp:168-176), an NCCL all-gather of the per-rank sorted (n,k) tables
over NVLink, and the device merge kernel (csrc/select.cu). Each (query, probe) pair lands on exactly
one rank, so total scan bytes do not grow with the number of GPUs.

Coarse quantization is the common step of IndexShardsIVF::search (faiss/IndexShardsIVF.cpp:175-183:
ONE quantizer->search, then per-shard search_preassigned). Here every rank holds the full centroid
table, so the common step is split by QUERY: rank r quantizes queries [r*n/R, (r+1)*n/R), the
[n][nprobe] probe table is all-gathered (n*nprobe*8 bytes in total) and every rank runs
search_preassigned on all the queries against its own lists. Collectives per search: that
all-gather and the (n*k*12 bytes per rank) result all-gather.
"""
from .binding import IndexIVFFlat, merge_knn_results


def list_range_of_rank(nlist, rank, world):
    """Lists [lo, hi) owned by `rank`: same arithmetic as faiss/gpu/GpuCloner.cpp:304-305."""
    return (rank * nlist) // world, ((rank + 1) * nlist) // world


def gather_results(D_loc, I_loc, group=None, out=None):
    """All-gather the per-rank (n,k) tables into [world][n][k] tensors (NCCL on GPU, gloo on CPU).
    out: optional (all_D, all_I) scratch of shape [world*n][k] to gather into."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    n, k = D_loc.shape
    # rank r's table lands in rows [r*n, (r+1)*n): the [world][n][k] layout merge_knn_results reads
    if out is not None:
        all_D, all_I = out
    else:
        all_D = torch.empty((world * n, k), dtype=D_loc.dtype, device=D_loc.device)
        all_I = torch.empty((world * n, k), dtype=I_loc.dtype, device=I_loc.device)
    dist.all_gather_into_tensor(all_D, D_loc.contiguous(), group=group)
    dist.all_gather_into_tensor(all_I, I_loc.contiguous(), group=group)
    return all_D.view(world, n, k), all_I.view(world, n, k)


class IndexShardsIVFFlat:
    """One rank's view of a list-sharded IndexIVFFlat.

    Every rank calls every method with the same arguments (SPMD). ``local`` may be injected (tests
    use a CPU stand-in over gloo); by default it is the device index restricted to this rank's
    list range.
    """

    def __init__(self, d, nlist, metric, device=0, group=None, local=None, merge=None,
                 shard_quantizer=True):
        import torch.distributed as dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.d, self.nlist, self.metric, self.device = d, nlist, metric, device
        self.list_lo, self.list_hi = list_range_of_rank(nlist, self.rank, self.world)
        if local is None:
            local = IndexIVFFlat(d, nlist, metric, device)
        local.set_list_range(self.list_lo, self.list_hi)
        self.local = local
        self.shard_quantizer = shard_quantizer
        self._merge = merge or (lambda m, aD, aI: merge_knn_results(m, aD, aI, device=device))

    @property
    def nprobe(self):
        return self.local.nprobe

    @nprobe.setter
    def nprobe(self, v):
        self.local.nprobe = v

    def set_centroids(self, c):
        self.local.set_centroids(c)

    def add_with_ids(self, x, ids):
        """Every rank is shown the same vectors and keeps those assigned to its lists."""
        self.local.add_with_ids(x, ids)

    def add_core(self, x, ids, assign):
        self.local.add_core(x, ids, assign)

    def search(self, x, k, broadcast_from=None):
        """x: [n][d] on this rank's device (replicated, or broadcast from rank `broadcast_from`).
        With ``self.trace = True`` (CUDA only) the device time of each phase of the call is left in
        ``self.last_phase_ms`` (broadcast, quantize + probe all-gather, local search, result
        all-gather, merge): one synchronize per call, so not for timed runs."""
        marks = []

        def mark(name):
            if getattr(self, "trace", False):
                import torch
                e = torch.cuda.Event(enable_timing=True)
                e.record()
                marks.append((name, e))

        mark("start")
        if broadcast_from is not None:
            import torch.distributed as dist
            dist.broadcast(x, src=broadcast_from, group=self.group)
        mark("broadcast")
        if self.world == 1:
            return self.local.search(x, k)
        if self.shard_quantizer:
            Iq = self._quantize_sharded(x)
            mark("quantize+gather")
            D_loc, I_loc = self.local.search_preassigned(x, k, Iq)
        else:
            D_loc, I_loc = self.local.search(x, k)
        mark("local search")
        import torch
        D_loc, I_loc = torch.as_tensor(D_loc), torch.as_tensor(I_loc)
        n_, k_ = D_loc.shape
        out = (self._scratch("all_D", (self.world * n_, k_), D_loc.dtype, D_loc.device),
               self._scratch("all_I", (self.world * n_, k_), I_loc.dtype, I_loc.device))
        all_D, all_I = gather_results(D_loc, I_loc, self.group, out=out)
        mark("result gather")
        out = self._merge(self.metric, all_D, all_I)
        mark("merge")
        if marks:
            import torch
            torch.cuda.synchronize()
            self.last_phase_ms = {marks[i][0]: marks[i - 1][1].elapsed_time(marks[i][1])
                                  for i in range(1, len(marks))}
        return out

    def _scratch(self, name, shape, dtype, device):
        """per-shape scratch tensors kept across calls (no allocator traffic in the search loop)"""
        import torch
        cache = self.__dict__.setdefault("_bufs", {})
        key = (name, tuple(shape), dtype, str(device))
        t = cache.get(key)
        if t is None:
            t = cache[key] = torch.empty(shape, dtype=dtype, device=device)
        return t

    def _quantize_sharded(self, x):
        """[n][nprobe] int64 probe table: each rank quantizes its slice of the queries straight
        into its rows of the gathered table, then one in-place all-gather. Probes of lists this
        rank does not own are simply skipped by its search_preassigned (the lists are empty there:
        faiss/IndexIVF.cpp:526-529)."""
        import torch
        import torch.distributed as dist
        n = x.shape[0]
        nprobe = min(self.local.nprobe, self.nlist)
        chunk = (n + self.world - 1) // self.world
        lo, hi = min(self.rank * chunk, n), min((self.rank + 1) * chunk, n)
        allq = self._scratch("allq", (self.world * chunk, nprobe), torch.int64, x.device)
        mine = allq[self.rank * chunk:(self.rank + 1) * chunk]
        if hi - lo < chunk:
            mine[hi - lo:].fill_(-1)
        if hi > lo:
            if _is_device_index(self.local) and x.is_cuda:
                dq = self._scratch("dq", (hi - lo, nprobe), torch.float32, x.device)
                self.local.quantize(x[lo:hi], nprobe, Dq=dq, Iq=mine[:hi - lo])
            else:
                _, Iq = self.local.quantize(x[lo:hi].contiguous(), nprobe)
                mine[:hi - lo] = torch.as_tensor(Iq).to(x.device)
        dist.all_gather_into_tensor(allq, mine, group=self.group)
        return allq[:n]


def _is_device_index(ix):
    return isinstance(ix, IndexIVFFlat)
