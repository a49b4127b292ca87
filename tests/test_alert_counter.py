"""the deterministic counter core of processFSEntry
(stream_process_alerts.js:348-434).  CPU: the C++ restatement against the statement-by-statement
Python twin on random fs entries.  GPU: the kernel fused behind the baseline against the oracle."""
import numpy as np
import pytest

from oracle import capi as ocapi, pyoracle as po

CAUSE_BITS = {c: 1 << i for i, c in enumerate(po.CAUSES)}


def _random_entries(rng, n):
    rec = np.zeros(n, dtype=ocapi.FS_DTYPE)
    rec["flags"] = 1
    rec["tpm"] = np.round(rng.choice([0.0, 0.6, 1.0, 1.2, 40.0], n), 2)
    rec["val"][:, 0] = np.round(rng.choice([np.nan, 50.0, 199.9, 200.0, 200.1, 900.0, 12000.0], n), 1)
    rec["val"][:, 1] = np.round(rng.choice([np.nan, 80.0, 250.5, 3000.0, 10000.0, 10000.5], n), 1)
    rec["signal"][:, 0] = rng.choice([-1, 0, 1, 1], n)
    rec["signal"][:, 1] = rng.choice([-1, 0, 1, 1], n)
    return rec


@pytest.mark.parametrize("both_only,window,required", [(True, 60, 45), (False, 60, 45), (True, 5, 3), (False, 1, 1), (True, 0, 0)])
def test_cpp_counter_matches_python_twin(both_only, window, required):
    rng = np.random.default_rng(window * 7 + required + both_only)
    rec = _random_entries(rng, 4000)
    rules = ocapi.make_alert_rules(window=window, required=required, alert_on_both_only=both_only)
    twin = po.AlertsCounter({"alertOnBothOnly": both_only, "rollingAlertWindowSizeInIntervals": window,
                             "requiredNumberBadIntervalsInAlertWindowToTrigger": required,
                             "overrides": {"services": {"svcB": {"hardMaxMsAlertThreshold": 2500}}}})
    counts = {}
    triggers = 0
    for i, r in enumerate(rec):
        service = ("svcA", "svcB", "S:getSalesServiceOffers")[i % 3]
        lag = ("360", "8640")[(i // 3) % 2]
        en = po.FullStatEntry(1, "srv", service, float(r["tpm"]), lag, r["val"][0], None, None, None, int(r["signal"][0]),
                              r["val"][1], None, None, None, int(r["signal"][1]), float("nan"), None, None, None, 0)
        want_t, want_c = twin.processFSEntry(en)
        c0 = counts.get((service, lag), 0)
        got_t, got_bits, c1 = ocapi.process_fs_entry(rules, r, c0, service_suppressed=(service == "S:getSalesServiceOffers"),
                                                     hard_max_override=2500.0 if service == "svcB" else 0.0)
        counts[(service, lag)] = c1
        assert got_t == want_t and got_bits == sum(CAUSE_BITS[c] for c in want_c), i
        assert c1 == twin.recentAlertCounts["srv"][service][lag]
        triggers += got_t
    assert triggers > 0


@pytest.mark.gpu
def test_gpu_alert_counter_matches_oracle():
    from apmbackend_b200 import Engine
    from tests.gpu_util import assert_fs_equal, engine_config, oracle_config
    from tests.streams import make_stream
    from tests.test_oracle_twins import CASES
    c = dict(CASES[0], n=6000, span=160)
    cols = make_stream(51, c["n"], c["S"], c["J"], c["bucket_ms"], c["span"], shift_services=(0, 1, 2), elapsed_hi=60000)
    K = c["S"] * c["J"]
    kw = dict(window=6, required=4, alert_on_both_only=False, hard_max_ms=400.0, hard_min_ms=30.0, hard_min_tpm=0.5)
    eng = Engine(engine_config(c))
    eng.set_alert_rules(suppressed_lags=(c["lags"][1],), **kw)
    hard = np.zeros(K); hard[2] = 150.0
    sup = np.zeros(K, np.uint8); sup[4] = 1
    eng.set_alert_key_params(0, hard, sup)
    orc = ocapi.Pipeline(oracle_config(c), fast=True)
    orc.set_alert_rules(ocapi.make_alert_rules(suppressed_lag_mask=0b10, **kw))
    orc.set_alert_key_params(2, 150.0, False)
    orc.set_alert_key_params(4, 0.0, True)
    orc.consume(*cols)
    eng.process_host(*cols)
    fs, al = eng.fetch_fs(), eng.fetch_alerts()
    want = orc.fs()
    assert_fs_equal(fs, want)  # includes flags (triggerAlert)
    key = lambda a: a[np.lexsort((a["lag"], a["key"], a["tick_ts"]))]  # noqa: E731
    assert np.array_equal(key(fs)["pad"], key(want)["pad"])  # causes and bad-interval counts
    trig = want[(want["flags"] & 2) != 0]
    assert trig.size > 0 and al.size == trig.size  # the alert block now holds the true alerts only
    assert key(al).tobytes() == key(trig).tobytes()
    assert not np.any(trig["lag"] == c["lags"][1]) and not np.any(trig["key"] == 4)
    eng.close()
