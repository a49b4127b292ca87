"""Host-side mirror of the reference's entry types (no device needed): CSV formatting
and parsing must equal the Python twin of entries.js byte for byte."""
import math

import numpy as np

from apmbackend_b200.binary_heap import BinaryHeap
from apmbackend_b200.entries import EntryFactory, FullStatEntry, StatEntry, TxEntry, to_fixed
from oracle import pyoracle as po


def test_to_fixed_equals_twin():
    rng = np.random.default_rng(5)
    xs = np.concatenate([rng.uniform(-500, 70000, 4000), rng.integers(0, 280000, 2000) / 4.0,
                         rng.integers(0, 1400000, 2000) / 20.0])
    for x in xs:
        for d in (0, 1, 2):
            assert to_fixed(float(x), d) == po.js_to_fixed(float(x), d)


def test_stat_and_fs_lines_equal_twin():
    nan = float("nan")
    for args in [(1578409230000, "xcla9571", "S:getFoo", 1.2, 141.66666666666666, 250.0, 300.0),
                 (1578409140000, "a", "b", 0.0, nan, nan, nan), (5, "a", "b", 8.345, 0.25, 0.35, 1234.05)]:
        assert StatEntry(*args).toCSVString() == po.StatEntry(*args).to_csv()
    fs_args = (1578409230000, "srv", "svc", 1.2, "360", 141.7, 100.04, -0.04, 300.25, 1, 250.0, None, None, None, 0,
               300.0, 99.95, 50.0, 150.0, -1)
    assert FullStatEntry(*fs_args).toCSVString() == po.FullStatEntry(*fs_args).to_csv()


def test_csv_round_trip_through_factory():
    f = EntryFactory()
    tx = f.getEntryFromCSV("tx|xcla9571|S:getFoo|abc123|42|1578409201000|1578409201120|120|true")
    assert isinstance(tx, TxEntry) and tx.endTs == 1578409201120 and tx.elapsed == 120
    assert tx.toCSVString() == "tx|xcla9571|S:getFoo|abc123|42|1578409201000|1578409201120|120|true"
    st = f.getEntryFromCSV("st|1578409230000|xcla9571|S:getFoo|1.20|141.7|250.0|undefined")
    assert st.average == 141.7 and math.isnan(st.per95)
    assert st.toCSVString() == "st|1578409230000|xcla9571|S:getFoo|1.20|141.7|250.0|undefined"
    line = ("fs|1579895320000|undefined|reportAbilityToPay|6|0.00|undefined:undefined:undefined:undefined:0|"
            "undefined:undefined:undefined:undefined:0.0|undefined:undefined:undefined:undefined:0.0")
    assert f.getEntryFromCSV(line).toCSVString() == line  # stream_process_alerts.js:638


def test_binary_heap_orders_by_score():
    h = BinaryHeap(lambda e: e["endTs"])
    for t in (5, 3, 9, 3, 1, 7):
        h.push({"endTs": t})
    assert h.size() == 6 and h.peek()["endTs"] == 1
    assert [e["endTs"] for e in h.popAllLessOrEqualToScore(5)] == [1, 3, 3, 5]
    assert h.size() == 2 and h.pop()["endTs"] == 7
