#!/usr/bin/env python
"""tools/shard_phases.py -- per-phase device time of the list-sharded search (torchrun, N ranks):
C2-shaped index (2M x 128 IP, nlist 4096, nprobe 24, 10k queries) split over the ranks."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])
import hnsw_ivf_b200 as b2  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
d, nb, nlist, nprobe, k, nq = 128, 2_000_000, 4096, 24, 10, 10_000
g = torch.Generator(device=dev)
g.manual_seed(1)
xb = torch.randn((nb, d), generator=g, device=dev)
xb /= xb.norm(dim=1, keepdim=True)
xq = torch.randn((nq, d), generator=g, device=dev)
xq /= xq.norm(dim=1, keepdim=True)
cent = xb[:nlist].contiguous()
ix = b2.IndexShardsIVFFlat(d, nlist, 0, device=local)
ix.set_stream(torch.cuda.current_stream().cuda_stream)
ix.set_centroids(cent)
lo, hi = rank * nb // world, (rank + 1) * nb // world
ix.add_with_ids(xb[lo:hi].contiguous(), torch.arange(lo, hi, device=dev, dtype=torch.int64))
ix.nprobe = nprobe
for _ in range(5):
    ix.search(xq, k, broadcast_from=0)
# every rank brings 1/world of the vectors (distributed add); phase times come from the C++
# object's CUDA events (b200_IndexShardsIVFFlat_phase_time_sums)
torch.cuda.synchronize()
dist.barrier()
ix.reset_phase_times()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    ix.search(xq, k, broadcast_from=0)
e1.record()
torch.cuda.synchronize()
calls, acc = ix.phase_time_sums()
if rank == 0:
    print("world", world, "ms per step", round(e0.elapsed_time(e1) / 20, 3),
          {n: round(v / max(calls, 1), 3) for n, v in acc.items()},
          "lib", {n: round(v, 3) for n, v in ix.local.last_timings().items()})
dist.destroy_process_group()
