import os, sys, torch
sys.path.insert(0, '/root/repo')
import hnsw_ivf_b200 as b2
d, nlist, n = 128, 4096, 10000
torch.manual_seed(0)
cent = torch.randn(nlist, d, device='cuda'); xq = torch.randn(n, d, device='cuda')
g = b2.IndexIVFFlat(d, nlist, 0, device=0)
g.set_centroids(cent)
for _ in range(2): g.quantize(xq, int(os.environ.get("KC", 1)))
torch.cuda.synchronize()
