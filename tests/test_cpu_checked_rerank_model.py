"""Property test of the CLAIM behind the checked re-rank (csrc/select.cu rerank_proven, DESIGN.md
section 3), on a small numpy model of the pipeline -- no GPU, no library code:

  scan      every (query, list) item keeps its KP best rows by APPROXIMATE score (|approx - exact| <= E)
            and everything it drops scored at least theta = min over full items of their KP-th score
  select    merge the kept keys by approximate score; A = k-th smallest; re-score exactly every key
            with approx <= A + 2 E among the first kk; B = smallest approximate score NOT re-scored
            (next merged key, theta); "proven" iff B - E > exact k-th score of the re-scored set

  claim     proven  =>  the k best re-scored keys ARE the exact top k of all rows, ties included
            (keys compare as (score, row), the library's order)

The model draws scores with gaps far below, around and far above E, so that both verdicts occur;
a wrong "proven" anywhere fails the test."""
import numpy as np


def model_search(exact, approx, lists, k, KP, kk, E):
    """returns (proven, result keys [(score, row)]) following select.cu"""
    kept, theta = [], np.inf
    for rows in lists:
        order = sorted(rows, key=lambda r: (approx[r], r))
        kept += order[:KP]
        if len(order) >= KP:
            theta = min(theta, approx[order[KP - 1]])   # what the item published (qthr)
    merged = sorted(kept, key=lambda r: (approx[r], r))
    if len(merged) < k:
        # fewer than k keys anywhere: proven only if nothing was dropped at all
        res = sorted(((exact[r], r) for r in merged))
        return theta == np.inf, res
    A = approx[merged[k - 1]]
    head = merged[:kk]
    nre = max(k, sum(1 for r in head if approx[r] <= A + 2 * E))
    rescored = sorted((exact[r], r) for r in merged[:nre])
    B = min(approx[merged[nre]] if len(merged) > nre else np.inf, theta)
    s_k = rescored[k - 1][0]
    return (B - E > s_k) or B == np.inf, rescored[:k]


def test_proven_results_are_exact():
    rng = np.random.default_rng(11)
    n_proven = n_unproven = 0
    for trial in range(3000):
        nlists = int(rng.integers(1, 9))
        n = int(rng.integers(5, 400))
        k = int(rng.integers(1, 12))
        KP = k + int(rng.integers(0, 8))
        kk = k + int(rng.integers(0, KP - k + 1))
        E = 10.0 ** rng.uniform(-4, -1)
        # scores: clusters of near-equal values (gaps << E), isolated ones (gaps >> E), exact ties
        base = rng.uniform(0, 1, n)
        if trial % 3 == 0:
            base = np.round(base, 2)                       # many exact ties
        elif trial % 3 == 1:
            base = np.round(base, 1) + rng.uniform(-E, E, n) * rng.uniform(0, 3)
        exact = base.astype(np.float64)
        approx = exact + rng.uniform(-E, E, n)             # |approx - exact| <= E
        if trial % 5 == 0:                                 # adversarial: errors at the extremes
            approx = exact + rng.choice([-E, E], n)
        owner = rng.integers(0, nlists, n)
        lists = [list(np.nonzero(owner == l)[0]) for l in range(nlists)]
        proven, res = model_search(exact, approx, lists, k, KP, kk, E)
        truth = sorted((exact[r], r) for r in range(n))[:k]
        if proven:
            n_proven += 1
            assert res == truth, (trial, k, KP, kk, E)
        else:
            n_unproven += 1
    # both verdicts must have been exercised for the test to mean anything
    assert n_proven > 300 and n_unproven > 300, (n_proven, n_unproven)


def test_wide_gaps_are_always_proven():
    """rows whose scores are more than 2 E apart: the window holds k keys and the proof goes through
    whenever KP leaves one spare key per item"""
    rng = np.random.default_rng(12)
    for _ in range(200):
        n, k, E = 200, 10, 1e-3
        exact = np.sort(rng.permutation(n) * 0.01 + rng.uniform(0, 1e-3, n))
        rng.shuffle(exact)
        approx = exact + rng.uniform(-E, E, n)
        owner = rng.integers(0, 4, n)
        lists = [list(np.nonzero(owner == l)[0]) for l in range(4)]
        proven, res = model_search(exact, approx, lists, k, KP=16, kk=16, E=E)
        assert proven
        assert res == sorted((exact[r], r) for r in range(n))[:k]
