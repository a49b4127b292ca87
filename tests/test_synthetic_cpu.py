"""The synthetic stream's popularity table (bench.py workloads): a monotone 32-bit CDF, the two-tier skew of config 5."""
import numpy as np

from apmbackend_b200.synthetic import zipf_cdf


def _probs(cdf):
    c = np.concatenate([[0.0], cdf.astype(np.float64) + 1.0]) / 4294967296.0
    return np.diff(c)


def test_zipf_cdf_is_monotone_and_complete():
    for k, s in ((1, 1.0), (2, 1.0), (10_000, 1.0), (400, 1.1)):
        cdf = zipf_cdf(k, s)
        assert cdf.dtype == np.uint32 and cdf.size == k and cdf[-1] == 0xFFFFFFFF
        assert np.all(np.diff(cdf.astype(np.int64)) >= 0)
    p = _probs(zipf_cdf(1000, 1.0))
    assert abs(p[0] / p[9] - 10.0) < 0.01  # 1 / rank


def test_two_tier_skew_gives_the_hot_services_their_share():
    p = _probs(zipf_cdf(10_000, 1.0, hot=(10, 0.8)))
    assert abs(p[:10].sum() - 0.8) < 1e-6 and np.allclose(p[:10], 0.08, atol=1e-6)
    assert abs(p[10:].sum() - 0.2) < 1e-6
    # every service is hot: no tail left to rescale (a shard that holds nothing but hot services)
    p = _probs(zipf_cdf(2, 1.0, hot=(2, 1.0)))
    assert np.allclose(p, 0.5, atol=1e-6)
