"""bench.py's multi-GPU job lets every rank generate its own key shard of ONE global synthetic stream on the device
(owner filter inside the generator).  That generator must produce exactly what binning the whole stream on the host
does - the same transactions in the same order with the same dense ids, and the global stream's ticks as marks -
or the sharded bench would not be the job it says it is."""
import numpy as np
import pytest

from apmbackend_b200 import sharding
from apmbackend_b200 import synthetic as syn

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("nranks,table", [(2, False), (3, True), (8, False)])
def test_device_shards_equal_the_host_partition_of_the_global_stream(nranks, table):
    spec = syn.StreamSpec(seed=9, n=150_000, num_services=300, num_jvms=2, span_ms=25 * 10_000, jitter_ms=1500, shift_mod=7)
    dev, host = syn.SynthStream(spec, device=True), syn.SynthStream(spec, device=False)
    svc, jvm, ts, el = host.generate_host()
    K = spec.num_services * spec.num_jvms
    owner = sharding.balanced_owner_table(syn.expected_key_counts(host), nranks, key_cost=50) if table else None
    plan = syn.ShardPlan(K, nranks, owner)
    latest = int(ts.min() // 10_000)  # as if an earlier step had ended in the stream's first bucket
    want, latest_out = sharding.partition_native(svc, jvm, ts, el, 10_000, spec.num_services, spec.num_jvms, nranks,
                                                 latest=latest, owner_table=owner, local_id=plan.local_id)
    total = 0
    for r in range(nranks):
        cols = syn.DeviceColumns(spec.n)
        n_r, marks, lat = syn.generate_shard_device(dev, plan, r, cols, 10_000, latest)
        got = [x[:n_r] for x in cols.to_host()]
        cols.free()
        assert lat == latest_out
        for a, b in zip(got, want[r][0]):
            assert np.array_equal(a, b)
        assert marks.tobytes() == want[r][1].tobytes()
        total += n_r
    assert total == spec.n
    # the expected per-key counts used to size the shards are the stream's real popularity
    key = jvm.astype(np.int64) * spec.num_services + svc
    exp = syn.expected_key_counts(host)
    real = np.bincount(key, minlength=K)
    assert abs(exp.sum() - spec.n) < 1 and np.corrcoef(exp, real)[0, 1] > 0.99
    plan.free(); dev.free()
