"""search() vs quantize() + search_preassigned() on one GPU, full index and one list-range shard."""
import sys, torch
sys.path.insert(0, '/root/repo')
import bench as B
import hnsw_ivf_b200 as b2
w = dict(B.WORKLOADS["C2"]); w["nb"] = 2_000_000; w["nlist"] = 1024
dev = torch.device("cuda", 0)
xb, xq, cent = B.setup_data(torch, w, dev)
d, nlist, metric, k, nprobe = w["d"], w["nlist"], w["metric"], 10, 32
for rng in (None, (0, nlist // 2)):
    g = b2.IndexIVFFlat(d, nlist, metric, device=0)
    g.set_centroids(cent)
    if rng:
        g.set_list_range(*rng)
    g.add_with_ids(xb, None)
    g.nprobe = nprobe
    D1, I1 = g.search(xq, k)
    _, Iq = g.quantize(xq, nprobe)
    # quantize in two halves like the sharded path
    h = xq.shape[0] // 2
    _, Ia = g.quantize(xq[:h].contiguous(), nprobe)
    _, Ib = g.quantize(xq[h:].contiguous(), nprobe)
    Iq2 = torch.cat([Ia, Ib])
    print("range", rng, "quantize halves == whole:", bool((Iq == Iq2).all()))
    D2, I2 = g.search_preassigned(xq, k, Iq2)
    print("  ids equal: %.6f  dis equal: %.6f" % ((I1 == I2).float().mean().item(), (D1 == D2).float().mean().item()))
    D3, I3 = g.search(xq, k)
    print("  search twice ids equal: %.6f" % ((I1 == I3).float().mean().item()))
