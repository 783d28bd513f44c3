import os, sys, torch, time
sys.path.insert(0, '/root/repo')
import hnsw_ivf_b200 as b2
d, nlist, n = int(os.environ.get("D", 128)), int(os.environ.get("NLIST", 4096)), 10000
torch.manual_seed(0)
cent = torch.randn(nlist, d, device='cuda'); xq = torch.randn(n, d, device='cuda')
g = b2.IndexIVFFlat(d, nlist, int(os.environ.get("METRIC", 0)), device=0)
g.set_centroids(cent)
g.set_stream(torch.cuda.current_stream().cuda_stream)
for kc in (1, 8, 32):
    for _ in range(3): g.quantize(xq, kc)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): g.quantize(xq, kc)
    e1.record(); torch.cuda.synchronize()
    print("S=%s kc=%d: %.3f ms" % (os.environ.get("B200IVF_COARSE_S"), kc, e0.elapsed_time(e1) / 20))
