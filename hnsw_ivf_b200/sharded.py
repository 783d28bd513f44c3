"""List-sharded IVF-Flat over several B200s, one process per GPU (torch.distributed / NCCL).

Behavioural model: IndexShardsIVF::search (faiss/IndexShardsIVF.cpp:158-243) -- a common coarse
quantizer, one search_preassigned per shard, merge_knn_results -- with the reference's host threads
and CPU merge replaced by: replicated device quantizer, contiguous list ranges per rank
(shard_type 4 / SUBSET_TYPE_INVLIST: faiss/gpu/GpuCloner.cpp:303-314,
faiss/invlists/InvertedLists.cpp:168-176), an NCCL all-gather of the per-rank sorted (n,k) tables
over NVLink, and the device merge kernel (csrc/select.cu). Each (query, probe) pair lands on exactly
one rank, so total scan bytes do not grow with the number of GPUs; the only collective is the
(n*k*12 byte per rank) result exchange.
"""
from .binding import IndexIVFFlat, merge_knn_results


def list_range_of_rank(nlist, rank, world):
    """Lists [lo, hi) owned by `rank`: same arithmetic as faiss/gpu/GpuCloner.cpp:304-305."""
    return (rank * nlist) // world, ((rank + 1) * nlist) // world


def gather_results(D_loc, I_loc, group=None):
    """All-gather the per-rank (n,k) tables into [world][n][k] tensors (NCCL on GPU, gloo on CPU)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    n, k = D_loc.shape
    # rank r's table lands in rows [r*n, (r+1)*n): the [world][n][k] layout merge_knn_results reads
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

    def __init__(self, d, nlist, metric, device=0, group=None, local=None, merge=None):
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
        """x: [n][d] on this rank's device (replicated, or broadcast from rank `broadcast_from`)."""
        if broadcast_from is not None:
            import torch.distributed as dist
            dist.broadcast(x, src=broadcast_from, group=self.group)
        D_loc, I_loc = self.local.search(x, k)
        if self.world == 1:
            return D_loc, I_loc
        all_D, all_I = gather_results(D_loc, I_loc, self.group)
        return self._merge(self.metric, all_D, all_I)
