"""Small seeded transaction streams shared by the CPU and GPU parity tests."""
import numpy as np


def make_stream(seed, n, num_services, num_jvms, bucket_ms, span_buckets, late_frac=0.01, late_ms=40000,
                shift_services=(), gap_at=None, t0=1578409200000, elapsed_hi=60000, sorted_time=True):
    """Returns (service_id u32, jvm_id u32, end_ts_ms i64, elapsed_ms i32) in arrival order.
    Zipf-ish service popularity, lognormal-ish durations, a fraction of late arrivals,
    optional 5x duration shift on some services in the last third (to fire signals),
    optional gap (no traffic) in the middle to exercise empty windows / bucket gaps."""
    rng = np.random.default_rng(seed)
    w = 1.0 / np.arange(1, num_services + 1) ** 1.1
    svc = rng.choice(num_services, size=n, p=w / w.sum()).astype(np.uint32)
    jvm = rng.integers(0, num_jvms, size=n).astype(np.uint32)
    t = np.sort(rng.integers(0, span_buckets * bucket_ms, size=n)) if sorted_time else \
        rng.integers(0, span_buckets * bucket_ms, size=n)
    if gap_at is not None:
        lo, hi = gap_at
        keep = (t < lo * bucket_ms) | (t >= hi * bucket_ms)
        svc, jvm, t = svc[keep], jvm[keep], t[keep]
        n = t.size
    mu = 3.0 + (svc % 7) * 0.4
    el = np.clip(np.rint(np.exp(rng.normal(mu, 0.5))), 1, elapsed_hi).astype(np.int64)
    for s in shift_services:
        m = (svc == s) & (t > (2 * span_buckets // 3) * bucket_ms)
        el[m] = np.minimum(el[m] * 5, elapsed_hi)
    late = rng.random(n) < late_frac
    t = t.astype(np.int64)
    t[late] -= rng.integers(0, late_ms, size=int(late.sum()))
    t = np.maximum(t, 0) + t0
    return svc, jvm, t.astype(np.int64), el.astype(np.int32)


def names(num_services, num_jvms):
    return [f"xcla{9500 + j}" for j in range(num_jvms)], [f"S:svc{i:04d}" for i in range(num_services)]
