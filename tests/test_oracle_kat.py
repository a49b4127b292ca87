"""Known-answer vectors of SURVEY.md section 8(c) against BOTH restatements
(oracle/apm_oracle.cpp via ctypes, oracle/pyoracle.py).  The reference has no
tests of its own (package.json:6-8), so these hand-derived vectors plus the
twin cross-check are what pins the oracle ("parity unpinned")."""
import math

import numpy as np
import pytest

from oracle import capi, pyoracle as po

BASE = 1578409200000
KAT1_TX = [(BASE + 1000, 120), (BASE + 2500, 80), (BASE + 9999, 200), (BASE + 10000, 100),
           (BASE + 15000, 300), (BASE + 30000, 50), (BASE + 100000, 70), (BASE + 110000, 90)]
KAT1_ST = [
    "st|1578409140000|xcla9571|S:getFoo|0.00|undefined|undefined|undefined",
    "st|1578409160000|xcla9571|S:getFoo|0.00|undefined|undefined|undefined",
    "st|1578409230000|xcla9571|S:getFoo|1.20|141.7|250.0|300.0",
    "st|1578409240000|xcla9571|S:getFoo|1.20|141.7|250.0|300.0",
]


def test_kat1_stats_stage_python_twin():
    st, fs, _, _ = po.run_pipeline([("xcla9571", "S:getFoo", t, e) for t, e in KAT1_TX])
    assert st == KAT1_ST
    assert len(fs) == 8  # two lags per st line


def test_kat1_stats_stage_cpp():
    cfg = capi.make_config(num_services=1, num_jvms=1)
    p = capi.Pipeline(cfg, fast=False)
    ts = np.array([t for t, _ in KAT1_TX], dtype=np.int64)
    el = np.array([e for _, e in KAT1_TX], dtype=np.int32)
    z = np.zeros(len(ts), dtype=np.uint32)
    p.consume(z, z, ts, el)
    lines = [capi.format_st(r, "xcla9571", "S:getFoo") for r in p.stats()]
    assert lines == KAT1_ST
    assert p.ticks == 5 and p.latest == (BASE + 110000) // 10000


KAT2 = {1: ((0,), (0,)), 2: ((1,), (1,)), 3: ((2,), (2,)), 4: ((2,), (3,)), 5: ((3, 4), (4,)), 8: ((5,), (7,)),
        10: ((7, 8), (9,)), 19: ((14, 15), (18,)), 20: ((14,), (18,)), 21: ((15, 16), (19, 20)),
        40: ((29,), (37,)), 100: ((74,), (94,)), 1000: ((749,), (949,))}


@pytest.mark.parametrize("n", sorted(KAT2))
def test_kat2_percentile_index_rule(n):
    a = [3 * i + 1 for i in range(n)]  # strictly increasing so indices are identifiable
    for p, idx in zip((75, 95), KAT2[n]):
        want = a[idx[0]] if len(idx) == 1 else (a[idx[0]] + a[idx[1]]) / 2
        assert po.calc_percentile(a, p) == want
        assert capi.calc_percentile(a, p) == want


def test_percentile_empty_and_edges():
    assert po.calc_percentile([], 75) is None
    assert math.isnan(capi.calc_percentile([], 75))
    assert capi.calc_percentile([5, 9], 0) == 5 and capi.calc_percentile([5, 9], 100) == 9
    for p in (1, 5, 10, 50, 75, 90, 95, 99, 99.9):
        for n in range(1, 130):
            a = list(range(10, 10 + n))
            assert capi.calc_percentile(a, p) == po.calc_percentile(a, p), (p, n)


def test_kat3_zscore():
    lst = [100.0] * 360
    for impl in (lambda *a: po.ZScoreParser().processZScoreStats(*a), capi.process_zscore):
        assert impl(360, 20.0, 0.1, 301.0, lst) == [120.1, 100.0, -100.0, 300.0, 1]
        assert impl(360, 20.0, 0.1, 300.0, lst) == [300.0, 100.0, -100.0, 300.0, 0]  # strict >
        r = impl(360, 20.0, 0.1, -100.1, lst)
        assert r[4] == -1 and r[0] == 0.1 * -100.1 + (1 - 0.1) * 100.0
        assert impl(360, 15.0, 0.0, 251.0, lst) == [100.0, 100.0, -50.0, 250.0, 1]
        # list shorter than lag: nothing defined, value pushed unchanged
        r = impl(360, 20.0, 0.1, 5.0, lst[:359])
        assert r[0] == 5.0 and r[4] == 0 and all(x is None or math.isnan(x) for x in r[1:4])
        # mean 0 -> "stddev" undefined -> no bounds, no signal
        r = impl(3, 20.0, 0.1, 5.0, [0.0, 0.0, 0.0])
        assert r[4] == 0 and r[1] == 0.0 and all(x is None or math.isnan(x) for x in r[2:4])
        # NaN new value never signals; NaN slots are skipped by average()
        r = impl(3, 1.0, 0.5, float("nan"), [4.0, float("nan"), 4.0])
        assert r[4] == 0 and r[1] == 4.0 and r[2] == 2.0 and r[3] == 6.0
        # last element NaN: signal fires, influence not applied
        r = impl(3, 1.0, 0.5, 100.0, [4.0, 4.0, float("nan")])
        assert r[4] == 1 and r[0] == 100.0


def test_stddev_is_sqrt_of_mean():
    """util_methods.js:41 passes an argument average() ignores."""
    lst = [1.0, 2.0, 3.0, 10.0]
    assert po.standard_deviation(lst) == math.sqrt(4.0) == capi.standard_deviation(lst)
    assert po.standard_deviation([]) is None and capi.standard_deviation([]) is None
    assert po.standard_deviation([0.0, 0.0]) is None and capi.standard_deviation([0.0, 0.0]) is None
    assert math.isnan(capi.standard_deviation([-4.0])) and math.isnan(po.standard_deviation([-4.0]))


KAT4 = [(33.33333333333333, 1, "33.3"), (0.05, 1, "0.1"), (0.15, 1, "0.1"), (0.25, 1, "0.3"), (0.35, 1, "0.3"),
        (1.45, 1, "1.4"), (1234.05, 1, "1234.0"), (8.345, 2, "8.35"), (0.0, 2, "0.00"), (-0.04, 1, "-0.0"),
        (-0.0, 1, "0.0"), (-100.0, 1, "-100.0"), (0.75, 1, "0.8"), (2.5, 0, "3"), (1e-30, 1, "0.0"),
        (123456789012.25, 1, "123456789012.3"), (0.125, 2, "0.13"), (0.375, 2, "0.38"), (1.005, 2, "1.00")]


@pytest.mark.parametrize("x,d,want", KAT4)
def test_kat4_to_fixed(x, d, want):
    assert po.js_to_fixed(x, d) == want
    assert capi.to_fixed(x, d) == want


def test_to_fixed_twins_agree_random():
    rng = np.random.default_rng(7)
    xs = np.concatenate([rng.uniform(0, 70000, 3000), rng.integers(0, 700000, 3000) / 20.0,
                         rng.integers(0, 70000, 500) / 4.0, -rng.uniform(0, 5000, 500),
                         rng.integers(1, 10**6, 2000) / rng.integers(1, 1000, 2000)])
    for x in xs:
        for d in (1, 2):
            s = po.js_to_fixed(float(x), d)
            assert capi.to_fixed(float(x), d) == s
            assert capi.quantize(float(x), d) == float(s)


def test_kat5_empty_window_fs_line():
    """The literal line at stream_process_alerts.js:638, '&' written as '|'."""
    want = ("fs|1579895320000|undefined|reportAbilityToPay|6|0.00|undefined:undefined:undefined:undefined:0|"
            "undefined:undefined:undefined:undefined:0.0|undefined:undefined:undefined:undefined:0.0")
    nan = float("nan")
    e = po.FullStatEntry(1579895320000, "undefined", "reportAbilityToPay", 0.0, 6, nan, None, None, None, 0,
                         nan, None, None, None, 0, nan, None, None, None, 0)
    assert e.to_csv() == want
    rec = np.zeros(1, dtype=capi.FS_DTYPE)
    rec["tick_ts"], rec["lag"], rec["tpm"] = 1579895320000, 6, 0.0
    for f in ("val", "avg", "lb", "ub"):
        rec[f] = nan
    assert capi.format_fs(rec[0], "undefined", "reportAbilityToPay") == want
