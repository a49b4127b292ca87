"""Stand-alone device operators against the oracle's util_methods.js restatement."""
import numpy as np
import pytest

from apmbackend_b200 import array_ops
from oracle import capi as ocapi

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n", [0, 1, 2, 4, 5, 20, 21, 100, 4095, 4096, 4097, 100_003])
def test_sorted_percentiles_match_calc_percentile(n):
    rng = np.random.default_rng(n)
    v = np.clip(np.rint(np.exp(rng.normal(4.5, 0.8, n))), 0, 600000).astype(np.int32)
    if n > 10:
        v[: n // 3] = v[0]  # heavy duplicates
    pcts = (75.0, 95.0, 50.0, 99.0)
    got, srt = array_ops.sorted_percentiles(v, pcts)
    want_sorted = np.sort(v, kind="stable")
    assert np.array_equal(srt, want_sorted)
    for g, p in zip(got, pcts):
        w = ocapi.calc_percentile(want_sorted, p)
        assert (np.isnan(g) and np.isnan(w)) or g == w, (n, p, g, w)


def test_sorted_percentiles_edges():
    got, srt = array_ops.sorted_percentiles([7, 7, 7], (0.0, 100.0))
    assert list(got) == [7.0, 7.0] and list(srt) == [7, 7, 7]
    big = np.array([2**31 - 1, 0, 65536, 65535], dtype=np.int32)
    got, srt = array_ops.sorted_percentiles(big, (75.0, 95.0))
    assert list(srt) == sorted(big.tolist())
    assert got[0] == ocapi.calc_percentile(np.sort(big), 75.0)
    with pytest.raises(Exception):
        array_ops.sorted_percentiles([1, -2, 3], (75.0,))
