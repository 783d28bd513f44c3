"""GPU tests at BASELINE.json's sizes, through size-independent properties (the oracle cannot run
these sizes in seconds): every kernel-selection setting gives the SAME result table (the exact
warp kernel is the oracle-checked one at small sizes, tests/test_gpu_parity.py), repeated searches
are identical, rows come out sorted, a list-range shard pair merges to the single index, and the
results of a query sample equal a brute force over the probed lists computed with torch.
Shapes: C1 (1M x 128 L2, nlist 1024, nprobe 16), C3 (1M x 960 L2: the 5-stage-ring shape whose
first cold search used to fail), a C2-shaped IP index (2M x 128, nlist 4096) for the mid-nprobe
regime where the two tensor-core kernels meet."""
import pytest

pytestmark = pytest.mark.gpu
L2, IP = 1, 0


def _data(torch, n, d, seed, normalize, ncl=4096):
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    cg = torch.Generator(device="cuda")
    cg.manual_seed(7)
    centers = torch.randn((ncl, d), generator=cg, device="cuda")
    which = torch.randint(0, ncl, (n,), generator=g, device="cuda")
    x = centers[which] + 0.35 * torch.randn((n, d), generator=g, device="cuda")
    if normalize:
        x /= x.norm(dim=1, keepdim=True)
    x = x.contiguous()
    # the handles below run on their own (non-blocking) streams, which are not ordered with torch's
    torch.cuda.synchronize()
    return x


def _build(torch, b2, d, nb, nlist, metric):
    xb = _data(torch, nb, d, 1, metric == IP)
    g = torch.Generator(device="cuda")
    g.manual_seed(3)
    cent = xb[torch.randperm(nb, generator=g, device="cuda")[:nlist]].contiguous()
    torch.cuda.synchronize()
    ix = b2.IndexIVFFlat(d, nlist, metric)
    ix.set_centroids(cent)
    ix.add(xb)
    return xb, cent, ix


def _same_tables(Da, Ia, Db, Ib, what):
    """two runs of the path: identical distances, ids identical except inside exact-tie groups"""
    assert (Da == Db).all().item(), what + ": distances differ"
    neq = Ia != Ib
    if neq.any().item():
        q, j = neq.nonzero(as_tuple=True)
        # a differing id must sit in a run of equal distances
        left = (Da[q, j] == Da[q, (j - 1).clamp(min=0)]) & (j > 0)
        right = (Da[q, j] == Da[q, (j + 1).clamp(max=Da.shape[1] - 1)]) & (j < Da.shape[1] - 1)
        assert (left | right).all().item(), what + ": ids differ outside tie groups"


@pytest.mark.parametrize("name,d,nb,nlist,metric,nprobes", [
    ("C1", 128, 1_000_000, 1024, L2, (16,)),
    ("C3", 960, 1_000_000, 1024, L2, (1, 16)),
    ("C2-shaped", 128, 2_000_000, 4096, IP, (4, 32, 64)),
])
def test_kernel_choices_agree_at_full_size(cuda_device, name, d, nb, nlist, metric, nprobes):
    import torch
    import hnsw_ivf_b200 as b2
    xb, cent, ix = _build(torch, b2, d, nb, nlist, metric)
    nq, k = 10_000, 10
    xq = _data(torch, nq, d, 2, metric == IP)
    a = torch.empty(nb, dtype=torch.int64, device="cuda")  # list of every vector, by torch
    step = 1 << 16
    for i0 in range(0, nb, step):
        xs = xb[i0:i0 + step]
        a[i0:i0 + step] = torch.cdist(xs, cent).argmin(1) if metric == L2 else (xs @ cent.T).argmax(1)
    for nprobe in nprobes:
        ix.nprobe = nprobe
        for name_, v in (("use_tc", 1), ("tc_mode", 0)):
            ix.set_option(name_, v)
        t0 = b2.tensor_core_launches()
        D0, I0 = ix.search(xq, k)  # the very first search of this handle: cold
        assert b2.tensor_core_launches() > t0, "expected a tensor-core scan at this shape"
        D1, I1 = ix.search(xq, k)
        assert (I0 == I1).all().item() and (D0 == D1).all().item(), "repeat differs"
        # sorted best first, padding only at the end
        if metric == L2:
            assert (D0[:, 1:] >= D0[:, :-1]).all().item()
        else:
            assert (D0[:, 1:] <= D0[:, :-1]).all().item()
        assert (I0 >= 0).all().item()
        # the other tensor-core kernel and the exact warp kernel give the same table
        for mode in (1, 2):
            ix.set_option("tc_mode", mode)
            Dm, Im = ix.search(xq, k)
            _same_tables(D0, I0, Dm, Im, "%s nprobe %d tc_mode %d" % (name, nprobe, mode))
        ix.set_option("tc_mode", 0)
        if d <= 128 or nprobe <= 1:  # (the exact kernel at d = 960, nprobe 16 takes seconds)
            ix.set_option("use_tc", 0)
            Ds, Is = ix.search(xq, k)
            _same_tables(D0, I0, Ds, Is, "%s nprobe %d exact kernel" % (name, nprobe))
            ix.set_option("use_tc", 1)
        # brute force over the probed lists for a query sample (torch, fp32 differences / products)
        _, Iq = ix.quantize(xq, nprobe)
        bad = 0
        for q in range(0, nq, nq // 12):
            rows = torch.isin(a, Iq[q]).nonzero().flatten()
            if metric == L2:
                dis = ((xb[rows] - xq[q]) ** 2).sum(1)
                top = torch.topk(dis, k, largest=False)
            else:
                dis = xb[rows] @ xq[q]
                top = torch.topk(dis, k, largest=True)
            want = set(rows[top.indices].tolist())
            got = set(I0[q].tolist())
            # (the device's own top-1 assignment may differ from cdist's on a near-tie: allow 1)
            bad += len(want ^ got) > 2
            torch.testing.assert_close(D0[q].sort().values, top.values.sort().values,
                                       rtol=1e-4, atol=1e-5)
        assert bad == 0


def test_shards_equal_single_at_c1_size(cuda_device):
    """two list-range shards on one device + device merge == the single index (the identity of
    tests/test_meta_index.py:128-176) at C1's size"""
    import torch
    import hnsw_ivf_b200 as b2
    d, nb, nlist, nprobe, k, nq = 128, 1_000_000, 1024, 16, 10, 10_000
    xb, cent, ix = _build(torch, b2, d, nb, nlist, L2)
    xq = _data(torch, nq, d, 2, False)
    ix.nprobe = nprobe
    D, I = ix.search(xq, k)
    _, Iq = ix.quantize(xq, nprobe)
    parts = []
    for r in range(2):
        s = b2.IndexIVFFlat(d, nlist, L2)
        s.set_centroids(cent)
        s.set_list_range(r * nlist // 2, (r + 1) * nlist // 2)
        s.add(xb)
        s.nprobe = nprobe
        parts.append(s.search_preassigned(xq, k, Iq))
        assert s.ntotal == nb
    allD = torch.stack([p[0] for p in parts]).contiguous()
    allI = torch.stack([p[1] for p in parts]).contiguous()
    torch.cuda.synchronize()
    Dm, Im = b2.merge_knn_results(L2, allD, allI)
    _same_tables(D, I, Dm, Im, "shards vs single")


def test_options_and_trap_note(cuda_device):
    import hnsw_ivf_b200 as b2
    from hnsw_ivf_b200.binding import debug_trap_info
    ix = b2.IndexIVFFlat(8, 4, L2)
    with pytest.raises(b2.B200Error) as e:
        ix.set_option("no_such_knob", 1)
    assert e.value.code == -2
    for name in ("simt_max_avg_q", "tcr_max_avg_q", "tc_mode", "use_tc", "use_graph"):
        ix.set_option(name, 1)
    assert debug_trap_info() is None  # no pipeline wait has ever given up in this process
