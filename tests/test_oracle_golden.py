"""Pins BOTH oracle back-ends (our C restatement and the reference's own code) against every
golden vector the reference's tests hold for this path:
  src/tests/rejection_test.c       GESDT / percentile / linear-fit on set1, set2
  src/tests/sorting.c              quickmedian == histogram_median == sorted median (u16)
  src/tests/stacking_blocks_test.c the 12 block-planner cases
"""
import json
import os

import numpy as np
import pytest

from oracle import oracle as O

G = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "rejection_sets.json")))
REJ = {"PERCENTILE": O.PERCENTILE, "LINEARFIT": O.LINEARFIT, "GESDT": O.GESDT}


@pytest.mark.parametrize("case", G["cases"], ids=[c["name"] for c in G["cases"]])
def test_rejection_golden(any_oracle, case):
    stack = np.array(G[case["set"]], dtype=np.float32)
    crit = None
    if case["rejection"] == "GESDT":
        crit = O.gesdt_critical_values(len(stack), *case["sig"])
        assert len(crit) == 7
    mean, lo, hi = any_oracle.mean_and_reject(stack, REJ[case["rejection"]], tuple(case["sig"]), crit=crit)
    assert [lo, hi] == case["rej"]
    # criterion's cr_expect_float_eq(actual, expected, eps) is an absolute tolerance on floats
    assert abs(np.float32(mean) - np.float32(case["mean"])) <= case["tol"] * max(1.0, abs(case["mean"]))


def test_gesdt_outlier_order(any_oracle):
    """rejection_test.c:89-93: outliers come out as 440, 410, 350, 3, 40 — equivalent to the kept
    set being everything except those five."""
    stack = np.array(G["set1"], dtype=np.float32)
    crit = O.gesdt_critical_values(22, 0.32, 0.05)
    mean, lo, hi = any_oracle.mean_and_reject(stack, O.GESDT, (0.32, 0.05), crit=crit)
    kept = [v for v in G["set1"] if v not in (440, 410, 350, 3, 40)]
    assert mean == pytest.approx(np.mean(np.array(kept, dtype=np.float64)), rel=1e-12)


@pytest.mark.parametrize("n", list(range(1, 41)) + [63, 64, 65, 100, 255, 256, 399, 400])
def test_median_property_u16(any_oracle, n):
    """src/tests/sorting.c:44-98"""
    rng = np.random.default_rng(n)
    a = rng.integers(0, 65536, size=n, dtype=np.uint16)
    srt = np.sort(a).astype(np.float64)
    expect = srt[n // 2] if n % 2 else (srt[n // 2 - 1] + srt[n // 2]) / 2.0
    m, _ = any_oracle.quickmedian(a)
    assert m == expect
    assert any_oracle.histogram_median(a) == expect


@pytest.mark.parametrize("n", [1, 2, 3, 5, 8, 9, 10, 16, 17, 100, 301])
def test_median_float(any_oracle, n):
    rng = np.random.default_rng(100 + n)
    a = rng.random(n, dtype=np.float32)
    srt = np.sort(a)
    if n % 2:
        expect = float(srt[n // 2])
    elif n < 9:   # sorting-network path averages in float (sorting.c:468-513)
        expect = float(np.float32(srt[n // 2 - 1] + srt[n // 2])) / 2.0
    else:
        expect = (float(srt[n // 2 - 1]) + float(srt[n // 2])) / 2.0
    m, _ = any_oracle.quickmedian(a)
    assert m == expect


@pytest.mark.parametrize("case", G["planner_cases"], ids=[f"test{i+1}" for i in range(len(G["planner_cases"]))])
def test_planner_golden(any_oracle, case):
    W, H, Cn = case["naxes"]
    rv, blocks, largest = any_oracle.planner(case["max_rows"], W, H, Cn, case["threads"])
    assert rv == 0
    nb = len(blocks)
    v = case["values"]
    assert {"in": nb in v, "eq": nb == v[0], "gt": nb > v[0], "ge": nb >= v[0]}[case["expect"]], nb
    # blocks tile the image (check_that_blocks_cover_the_image) and fit the budget
    y, z = 0, 0
    for ch, s, e, h in blocks:
        assert s == y and h == e - s + 1
        y = e + 1
        if y == H:
            y, z = 0, z + 1
    assert y == 0 and z == Cn
    assert largest * case["threads"] <= case["max_rows"]
    if "note" in case:   # the real-world report: 33 parallel blocks of size 365 (+9)
        assert nb == 33 and (H * Cn) // nb == 365
