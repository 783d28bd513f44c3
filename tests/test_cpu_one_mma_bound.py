"""The error bound the checked re-rank of the one-MMA filter mode relies on (csrc/index.cu,
kOneMmaErrCoef), checked numerically on the CPU with the tensor core's operand
handling emulated bit for bit: the query operand is rounded to tf32 (round to nearest, ties away:
add 0x1000, clear 13 bits -- what the kernel's prologue does), the row operand is TRUNCATED to tf32
(what tcgen05.mma kind::tf32 does with fp32 bit patterns, tools/tc_probe.cu), products are exact.

    |tf32_rn(x) . trunc_tf32(y) - x . y|  <=  (2^-11 + 2^-10 + 2^-21) |x| |y|

and the coefficient the library uses leaves 2^-13 |x||y| on top for every fp32 accumulation
involved. For L2 the library uses coef * (|x|^2 + |y|^2) >= 2 coef |x||y| against an error of
2 |dot error|."""
import re
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ROUNDING = 2.0 ** -11 + 2.0 ** -10 + 2.0 ** -21


def tf32_rn(a):
    u = a.astype(np.float32).view(np.uint32)
    return ((u + np.uint32(0x1000)) & np.uint32(0xFFFFE000)).view(np.float32)


def tf32_trunc(a):
    u = a.astype(np.float32).view(np.uint32)
    return (u & np.uint32(0xFFFFE000)).view(np.float32)


def coef_in_source():
    src = open(os.path.join(ROOT, "hnsw_ivf_b200", "csrc", "index.cu")).read()
    m = re.search(r"constexpr float kOneMmaErrCoef = ([^;]+);", src)
    assert m, "kOneMmaErrCoef not found"
    return float(eval(m.group(1).replace("f", "")))


def worst_ratio(x, y):
    """max over query / row pairs of |approx - exact| / (|x||y|), float64 arithmetic"""
    xh = tf32_rn(x).astype(np.float64)
    yh = tf32_trunc(y).astype(np.float64)
    approx = xh @ yh.T
    exact = x.astype(np.float64) @ y.astype(np.float64).T
    nx = np.linalg.norm(x.astype(np.float64), axis=1)[:, None]
    ny = np.linalg.norm(y.astype(np.float64), axis=1)[None, :]
    return float((np.abs(approx - exact) / (nx * ny)).max())


def test_library_coefficient_covers_the_operand_roundings():
    c = coef_in_source()
    assert c >= ROUNDING + 2.0 ** -13, (c, ROUNDING)
    assert c < 2.0 ** -9  # (a bound that wide would re-score far too much)


def test_bound_on_random_vectors():
    rng = np.random.default_rng(3)
    for d in (20, 96, 128, 960):
        x = (rng.standard_normal((64, d)) * rng.uniform(0.01, 100.0)).astype(np.float32)
        y = (rng.standard_normal((512, d)) * rng.uniform(0.01, 100.0)).astype(np.float32)
        assert worst_ratio(x, y) <= ROUNDING


def test_bound_on_adversarial_mantissas():
    """every element sits just below a tf32 step (truncation loses almost a whole step) and has the
    sign that makes the losses add up instead of cancelling"""
    rng = np.random.default_rng(4)
    d = 128
    base = rng.uniform(1.0, 2.0, (256, d)).astype(np.float32)
    y = (base.view(np.uint32) | np.uint32(0x1FFF)).view(np.float32)       # low 13 bits all ones
    x = (rng.uniform(1.0, 2.0, (32, d)).astype(np.float32).view(np.uint32) & np.uint32(0xFFFFE000)
         | np.uint32(0x0FFF)).view(np.float32)                            # just under a rounding tie
    r = worst_ratio(x, y)
    assert r <= ROUNDING
    assert r > 2.0 ** -11  # (the construction does get close to the bound: it is not slack)
    # mixed signs, wide dynamic range
    s = rng.choice([-1.0, 1.0], (256, d)).astype(np.float32)
    y2 = y * s * (2.0 ** rng.integers(-8, 8, (256, d))).astype(np.float32)
    assert worst_ratio(x, y2) <= ROUNDING


def test_l2_form_of_the_bound():
    """score = |y|^2 - 2 x.y + |x|^2: the library charges coef * (|x|^2 + |y|^2)"""
    rng = np.random.default_rng(5)
    c = coef_in_source()
    x = (rng.standard_normal((32, 96)) * 3).astype(np.float32)
    y = (rng.standard_normal((400, 96)) * 0.5 + 2.0).astype(np.float32)
    xh = tf32_rn(-2.0 * x).astype(np.float64)
    yh = tf32_trunc(y).astype(np.float64)
    xn = (x.astype(np.float64) ** 2).sum(1)[:, None]
    yn = (y.astype(np.float64) ** 2).sum(1)[None, :]
    approx = yn + xh @ yh.T + xn
    exact = ((x.astype(np.float64)[:, None, :] - y.astype(np.float64)[None, :, :]) ** 2).sum(2)
    assert (np.abs(approx - exact) <= c * (xn + yn)).all()
