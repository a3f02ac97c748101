"""The GESDT critical values need the Student-t quantile (gsl_cdf_tdist_Pinv in the reference,
median_and_mean.c:1482 — GSL is an un-vendored, unpinned system dependency).  The harness uses
scipy's t.ppf; this test pins it on a SECOND, independent implementation written here from the
textbook definitions — the t distribution function through the regularized incomplete beta function
(Lentz's continued fraction), inverted by bisection to full double precision — for every stack depth
3 .. 500 at alpha = 0.05 and 0.01: the float critical values must be equal, and the quantiles agree
to 1e-12.  Plus the reference's own golden answer (rejection_test.c: 25 samples, 0.32 / 0.05)."""
import math

import numpy as np

from oracle import oracle as O


def betacf(a, b, x):
    tiny = 1e-300
    qab, qap, qam = a + b, a + 1.0, a - 1.0
    c, d = 1.0, 1.0 - qab * x / qap
    d = tiny if abs(d) < tiny else d
    d = 1.0 / d
    h = d
    for m in range(1, 400):
        m2 = 2 * m
        aa = m * (b - m) * x / ((qam + m2) * (a + m2))
        d = 1.0 + aa * d; d = tiny if abs(d) < tiny else d
        c = 1.0 + aa / c; c = tiny if abs(c) < tiny else c
        d = 1.0 / d; h *= d * c
        aa = -(a + m) * (qab + m) * x / ((a + m2) * (qap + m2))
        d = 1.0 + aa * d; d = tiny if abs(d) < tiny else d
        c = 1.0 + aa / c; c = tiny if abs(c) < tiny else c
        d = 1.0 / d
        de = d * c
        h *= de
        if abs(de - 1.0) < 1e-16:
            break
    return h


def betai(a, b, x):
    if x <= 0.0:
        return 0.0
    if x >= 1.0:
        return 1.0
    bt = math.exp(math.lgamma(a + b) - math.lgamma(a) - math.lgamma(b) + a * math.log(x) + b * math.log1p(-x))
    if x < (a + 1.0) / (a + b + 2.0):
        return bt * betacf(a, b, x) / a
    return 1.0 - bt * betacf(b, a, 1.0 - x) / b


def t_cdf(t, nu):
    x = nu / (nu + t * t)
    tail = 0.5 * betai(0.5 * nu, 0.5, x)
    return 1.0 - tail if t > 0 else tail


def t_ppf(p, nu):
    lo, hi = 0.0, 1e6       # p > 0.5 here
    for _ in range(200):
        mid = 0.5 * (lo + hi)
        if t_cdf(mid, nu) < p:
            lo = mid
        else:
            hi = mid
        if hi - lo <= 1e-15 * hi:
            break
    return 0.5 * (lo + hi)


def critical_values_independent(nb_frames, sig0, sig1):
    f32 = np.float32
    n_out = int(np.floor(f32(nb_frames) * f32(sig0)))
    out = np.empty(n_out, dtype=np.float32)
    size = nb_frames
    for j in range(n_out):
        p = f32(1) - f32(sig1) / f32(2 * size)
        td = f32(t_ppf(float(p), size - 2))
        out[j] = f32(f32(size - 1) * td / (np.sqrt(f32(size)) * np.sqrt(f32(size - 2) + td * td)))
        size -= 1
    return out


def test_quantile_against_independent_implementation():
    from scipy.stats import t as tdist
    worst = 0.0
    for nu in list(range(1, 60)) + [75, 100, 150, 200, 298, 400, 498]:
        for p in (0.975, 0.99, 0.995, 0.9995, 0.99995, 1 - 0.05 / 1000):
            a, b = t_ppf(p, nu), float(tdist.ppf(p, nu))
            worst = max(worst, abs(a - b) / b)
    assert worst < 1e-11, worst


def test_critical_values_for_every_depth():
    for alpha in (0.05, 0.01):
        for n in list(range(3, 120)) + [150, 200, 256, 300, 400, 500]:
            a = O.gesdt_critical_values(n, 0.3, alpha)
            b = critical_values_independent(n, 0.3, alpha)
            assert a.shape == b.shape and np.array_equal(a, b), (n, alpha, a[:4], b[:4])


def test_count_is_a_float_product():
    """median_and_mean.c:1479: `(int) floor(nb_frames * sig[0])` multiplies an int by a FLOAT"""
    for n, s, want in ((10, 0.7, 7), (20, 0.35, 7), (40, 0.35, 14), (10, 0.9, 9), (300, 0.3, 90)):
        assert len(O.gesdt_critical_values(n, s, 0.05)) == want, (n, s)
