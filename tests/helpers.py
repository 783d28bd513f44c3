"""Shared test helpers: synthetic data and the tie-aware result comparison.

The comparison restates the reference's own tie-aware checker
(contrib/evaluation.py:243-262 check_ref_knn_with_draws) with a relative tolerance, because the
reference's distance kernels are compiled with reassociation (faiss/impl/platform_macros.h:176-181)
and no two builds agree to the last bit.
"""
import numpy as np

RTOL = 1e-4  # north_star: distances within 1e-4 relative


def clustered(n, d, ncenters, seed, sigma=0.35, normalize=False, centers_seed=77):
    """Gaussian mixture used by every config."""
    rc = np.random.default_rng(centers_seed)
    centers = rc.standard_normal((ncenters, d), dtype=np.float32)
    r = np.random.default_rng(seed)
    which = r.integers(0, ncenters, size=n)
    x = centers[which] + sigma * r.standard_normal((n, d), dtype=np.float32)
    if normalize:
        x /= np.linalg.norm(x, axis=1, keepdims=True)
    return np.ascontiguousarray(x, dtype=np.float32)


def sample_centroids(xb, nlist, seed=5):
    r = np.random.default_rng(seed)
    idx = r.choice(xb.shape[0], size=nlist, replace=False)
    return np.ascontiguousarray(xb[np.sort(idx)])


def brute_force(metric, xq, xb, k):
    """exact kNN in float64 (ground truth for recall)"""
    xq64, xb64 = xq.astype(np.float64), xb.astype(np.float64)
    if metric == 1:
        dis = (xq64 ** 2).sum(1)[:, None] + (xb64 ** 2).sum(1)[None, :] - 2 * xq64 @ xb64.T
        I = np.argsort(dis, axis=1, kind="stable")[:, :k]
    else:
        dis = xq64 @ xb64.T
        I = np.argsort(-dis, axis=1, kind="stable")[:, :k]
    return np.take_along_axis(dis, I, 1), I


def recall_at_k(I, I_gt):
    k = I_gt.shape[1]
    hits = sum(len(set(a.tolist()) & set(b.tolist())) for a, b in zip(I, I_gt))
    return hits / float(I_gt.shape[0] * k)


def _close(a, b, rtol, atol):
    return abs(a - b) <= rtol * max(abs(a), abs(b)) + atol


def compare_knn(D_ref, I_ref, D_new, I_new, rtol=RTOL, atol=1e-6):
    """Tie-aware comparison of two (n,k) result tables.

    Returns a dict with the exact position-wise id match rate and the number of UNEXPLAINED
    mismatches. A mismatch at (q, j) is explained when
      * the new id sits at another rank j' of the reference row whose reference distance is within
        tolerance of rank j's (a swap inside a group of (near-)equal distances), or
      * the new id is not in the reference row at all and its distance is within tolerance of the
        reference's last kept distance (a (near-)tie at the k-th boundary),
    and in both cases the distance at that rank agrees within tolerance.
    """
    D_ref, I_ref, D_new, I_new = map(np.asarray, (D_ref, I_ref, D_new, I_new))
    assert D_ref.shape == D_new.shape == I_ref.shape == I_new.shape
    n, k = I_ref.shape
    bad_dist = 0
    unexplained = 0
    exact = int((I_ref == I_new).sum())
    for q in range(n):
        neq = np.nonzero(I_ref[q] != I_new[q])[0]
        # distances must agree rank by rank (sorted tables; near-tie swaps keep this true)
        for j in range(k):
            if I_ref[q, j] == -1 and I_new[q, j] == -1:
                continue
            if not _close(float(D_ref[q, j]), float(D_new[q, j]), rtol, atol):
                bad_dist += 1
        if neq.size == 0:
            continue
        ref_pos = {int(i): j for j, i in enumerate(I_ref[q]) if i >= 0}
        last = max([j for j in range(k) if I_ref[q, j] >= 0], default=-1)
        for j in neq:
            nid = int(I_new[q, j])
            if nid in ref_pos:
                jp = ref_pos[nid]
                ok = _close(float(D_ref[q, jp]), float(D_ref[q, j]), 2 * rtol, atol)
            elif nid >= 0 and last >= 0:
                ok = _close(float(D_new[q, j]), float(D_ref[q, last]), 2 * rtol, atol)
            else:
                ok = False
            if not ok:
                unexplained += 1
    return {"exact_id_rate": exact / float(n * k), "unexplained": unexplained,
            "bad_dist": bad_dist, "n": n, "k": k}


def assert_knn_parity(D_ref, I_ref, D_new, I_new, rtol=RTOL, min_exact=0.995, what=""):
    r = compare_knn(D_ref, I_ref, D_new, I_new, rtol)
    assert r["bad_dist"] == 0, "%s: %d distances off by more than %g: %s" % (what, r["bad_dist"],
                                                                             rtol, r)
    assert r["unexplained"] == 0, "%s: %d id mismatches not explained by ties: %s" % (
        what, r["unexplained"], r)
    assert r["exact_id_rate"] >= min_exact, "%s: exact id rate %.5f" % (what, r["exact_id_rate"])
    return r


def compare_probes(Dq_ref, Iq_ref, Iq_new, Dq_new=None, rtol=RTOL, atol=1e-5):
    """Probe lists must be identical except where the reference's own coarse distances are
    (near-)tied: same rule as compare_knn applied to the coarse table."""
    D_new = Dq_new if Dq_new is not None else Dq_ref
    return compare_knn(Dq_ref, Iq_ref, D_new, Iq_new, rtol, atol)


def resolve_spec(spec, table):
    """Selector spec of a golden file (arrays given by the NAME of an entry of `table`) -> the spec
    oracle/pyoracle.py understands (arrays inline)."""
    if spec is None:
        return None
    kind = spec[0]
    if kind == "range":
        return ("range", int(spec[1]), int(spec[2]))
    if kind in ("batch", "bitmap"):
        a = spec[1]
        return (kind, np.asarray(table[a]) if isinstance(a, str) else np.asarray(a))
    if kind == "not":
        return ("not", resolve_spec(spec[1], table))
    return (kind, resolve_spec(spec[1], table), resolve_spec(spec[2], table))


def gpu_selector(spec, device=0):
    """Selector spec -> hnsw_ivf_b200 selector object (the C-ABI's IDSelector family)."""
    import hnsw_ivf_b200 as b2
    if spec is None:
        return None
    kind = spec[0]
    if kind == "range":
        return b2.IDSelectorRange(spec[1], spec[2])
    if kind == "batch":
        return b2.IDSelectorBatch(spec[1], device)
    if kind == "bitmap":
        return b2.IDSelectorBitmap(spec[1], device)
    if kind == "not":
        return b2.IDSelectorNot(gpu_selector(spec[1], device))
    ctor = {"and": b2.IDSelectorAnd, "or": b2.IDSelectorOr, "xor": b2.IDSelectorXOr}[kind]
    return ctor(gpu_selector(spec[1], device), gpu_selector(spec[2], device))


def compare_range(l_ref, D_ref, I_ref, l_new, D_new, I_new, radius, metric, rtol=RTOL):
    """Range results per query as sets: ids may differ only where the reference distance is within
    tolerance of the radius; common ids must agree on the distance. Returns the number of ids that
    differ (all explained by the boundary), raises on anything else."""
    n = len(l_ref) - 1
    assert len(l_new) - 1 == n
    boundary = 0
    for q in range(n):
        a = {int(i): float(v) for i, v in zip(I_ref[int(l_ref[q]):int(l_ref[q + 1])],
                                              D_ref[int(l_ref[q]):int(l_ref[q + 1])])}
        b = {int(i): float(v) for i, v in zip(I_new[int(l_new[q]):int(l_new[q + 1])],
                                              D_new[int(l_new[q]):int(l_new[q + 1])])}
        for i in set(a) & set(b):
            assert _close(a[i], b[i], rtol, 1e-6), (q, i, a[i], b[i])
        for i in set(a) ^ set(b):
            v = a.get(i, b.get(i))
            assert _close(v, radius, 4 * rtol, 1e-6), "query %d id %d dis %r radius %r" % (
                q, i, v, radius)
            boundary += 1
    return boundary
