"""b200_IndexIVFFlat_quantize_noalloc (no counterpart in the reference):
the coarse quantizer with its scratch in a fixed file-scope __device__ array, for
callers that cannot cudaMalloc. Checked against the existing entry point b200_IndexIVFFlat_quantize
through the C-ABI, on two calls in a row (the array is shared and keeps the last call's contents).
Collected last on purpose."""
import numpy as np
import pytest

from helpers import clustered, compare_knn

pytestmark = pytest.mark.gpu
L2, IP = 1, 0


@pytest.mark.parametrize("metric", [L2, IP])
def test_quantize_noalloc_matches_quantize_twice(cuda_device, metric):
    import torch
    import hnsw_ivf_b200 as b2
    d, nlist = 96, 3000
    cent = clustered(nlist, d, 256, 3, normalize=(metric == IP))
    g = b2.IndexIVFFlat(d, nlist, metric)
    g.set_centroids(cent)
    # two calls in a row with different shapes: the second reuses scratch the first left dirty
    for n, nprobe, seed in ((700, 16, 8), (333, 5, 9), (1, 64, 10)):
        xq = clustered(n, d, 256, seed, normalize=(metric == IP))
        xd = torch.from_numpy(xq).cuda()
        Dr, Ir = g.quantize(xd, nprobe)                      # existing entry point
        Dn = torch.empty((n, nprobe), dtype=torch.float32, device="cuda")
        In = torch.empty((n, nprobe), dtype=torch.int64, device="cuda")
        g.quantize_noalloc(xd, nprobe, Dn, In)               # new entry point (loads its kernels)
        torch.cuda.synchronize()
        first = (Dn.clone(), In.clone())
        Dn.fill_(-1.0)
        In.fill_(-7)
        free_before = torch.cuda.mem_get_info()[0]
        g.quantize_noalloc(xd, nprobe, Dn, In)               # ... and again, straight after
        torch.cuda.synchronize()
        # no device memory was taken, by the library or by anybody else, during the call
        assert torch.cuda.mem_get_info()[0] == free_before
        assert torch.equal(first[0], Dn) and torch.equal(first[1], In)
        Dr, Ir, Dn, In = (t.cpu().numpy() for t in (Dr, Ir, Dn, In))
        r = compare_knn(Dr, Ir, Dn, In, rtol=1e-4, atol=1e-5)   # the tolerance of the other tests
        assert r["bad_dist"] == 0 and r["unexplained"] == 0, r
        assert r["exact_id_rate"] >= 0.999, r
    # limits are refused with a clear message, host buffers too
    xd = torch.zeros((5000, d), dtype=torch.float32, device="cuda")
    Dn = torch.empty((5000, 4), dtype=torch.float32, device="cuda")
    In = torch.empty((5000, 4), dtype=torch.int64, device="cuda")
    with pytest.raises(b2.B200Error) as ei:
        g.quantize_noalloc(xd, 4, Dn, In)
    assert "exceeds the maximum" in str(ei.value)
    with pytest.raises(b2.B200Error) as ei:
        g.quantize_noalloc(xd[:10], 65, torch.empty((10, 65), device="cuda"),
                           torch.empty((10, 65), dtype=torch.int64, device="cuda"))
    assert "nprobe" in str(ei.value)
    with pytest.raises(b2.B200Error) as ei:
        g.quantize_noalloc(np.zeros((4, d), np.float32), 4, np.zeros((4, 4), np.float32),
                           np.zeros((4, 4), np.int64))
    assert "device buffers only" in str(ei.value)
