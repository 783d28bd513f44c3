#!/usr/bin/env python
"""tools/shard_phases.py -- per-phase device time of the list-sharded search (torchrun, N ranks):
C2-shaped index (2M x 128 IP, nlist 4096, nprobe 24, 10k queries) split over the ranks."""
import os
import statistics
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
ix.local.set_stream(torch.cuda.current_stream().cuda_stream)
ix.set_centroids(cent)
ix.add_with_ids(xb, None)
ix.nprobe = nprobe
for _ in range(5):
    ix.search(xq, k, broadcast_from=0)
ix.trace = True
acc = {}
for _ in range(20):
    ix.search(xq, k, broadcast_from=0)
    for name, v in ix.last_phase_ms.items():
        acc.setdefault(name, []).append(v)
ix.trace = False
torch.cuda.synchronize()
dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    ix.search(xq, k, broadcast_from=0)
e1.record()
torch.cuda.synchronize()
if rank == 0:
    print("world", world, "ms per step", round(e0.elapsed_time(e1) / 20, 3),
          {n: round(statistics.median(v), 3) for n, v in acc.items()},
          "lib", {n: round(v, 3) for n, v in ix.local.last_timings().items()})
dist.destroy_process_group()
