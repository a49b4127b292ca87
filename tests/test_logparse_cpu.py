"""Oracle of the log feeder (oracle/pylogparse.py) against known answers derived by hand from
stream_parse_transactions.js (no reference tests or fixtures exist for the parser: parity is unpinned)."""
import math

from oracle.pylogparse import LogParserOracle, TimeZone, convert_string_date_to_ms, js_parse_int

from tests.logs import make_log

FP = "/logs/jvm01/app.log"
SRV = "/logs/jvm02/server.log"


def run(fp, lines, tz=None):
    o = LogParserOracle(tz)
    for l in lines:
        o.readLine(fp, l)
    o.flush()
    return o.out


def test_convert_string_date_to_ms_known_answers():
    tz = TimeZone()
    # 2020-01-07T10:00:02.669Z = 1578391202669 (18268 days * 86400 s + 36002 s)
    assert convert_string_date_to_ms("2020-01-07 10:00:02,669", tz) == 1578391202669
    assert convert_string_date_to_ms("1970-01-01 00:00:00,000", tz) == 0
    assert convert_string_date_to_ms("2000-02-29 23:59:59,999", tz) == 951868799999
    assert convert_string_date_to_ms("", tz) == ""
    assert math.isnan(convert_string_date_to_ms("undefined undefined", tz))
    # the Date constructor carries out-of-range fields: month 13 of 2019 is January 2020, day 0 the last of the month before
    assert convert_string_date_to_ms("2019-13-07 10:00:02,669", tz) == 1578391202669
    assert convert_string_date_to_ms("2020-01-00 24:00:00,000", tz) == convert_string_date_to_ms("2020-01-01 00:00:00,000", tz)
    # US Central: UTC-6 in winter, UTC-5 after the second Sunday of March 02:00 local
    cst = TimeZone([(-(1 << 62), -6 * 3600000), (1583632800000, -5 * 3600000)])  # 2020-03-08 02:00 local
    assert convert_string_date_to_ms("2020-01-07 04:00:02,669", cst) == 1578391202669
    assert convert_string_date_to_ms("2020-03-08 03:00:00,000", cst) == 1583654400000  # 08:00Z


def test_parse_int():
    assert js_parse_int("569") == 569 and js_parse_int("12ms") == 12 and js_parse_int(" 7") == 7 and js_parse_int("-3") == -3
    assert math.isnan(js_parse_int("time")) and math.isnan(js_parse_int("")) and math.isnan(js_parse_int("undefined"))


def test_start_stop_pair_goes_through_the_need_number_cache():
    out = run(FP, ["[abc123] 2020-01-07 10:00:02,100 [t1] INFO  CommonTiming::Start Provider[acct].lookup",
                   "[abc123] 2020-01-07 10:00:02,669 [t1] INFO  CommonTiming::Stop Provider[acct].lookup elapsed time = 569 ms"])
    # `[t1] INFO ` matches the BAF pattern; "t1" is not a number, so the account number stays NaN (:296)
    assert out == ["tx|jvm01|Provider:acct.lookup|abc123|NaN|1578391202100|1578391202669|569|N"]


def test_stop_without_start_is_salvaged_with_an_empty_log_id():
    out = run(FP, ["[zz9] 2020-01-07 10:00:03,000 INFO  CommonTiming::Stop Svc elapsed time = 5 ms"])
    assert out == ["tx|jvm01|Svc||NaN|1578391202995|1578391203000|5|N"]  # start = end - elapsed (:279-281)


def test_empty_log_id():
    out = run(FP, ["[] 2020-01-07 10:00:02,100 INFO  CommonTiming::Start Svc",
                   "[] 2020-01-07 10:00:03,000 INFO  CommonTiming::Stop Svc elapsed time = 900 ms"])
    assert out == ["tx|jvm01|Svc||NaN|1578391202100|1578391203000|900|N"]  # the Start was discarded (:466): 3000 - 900


def test_baf_account_number_completes_the_record_at_once():
    out = run(FP, ["[q1] 2020-01-07 10:00:02,000 [thr][sess:4455667] INFO  CommonTiming::Start Svc",
                   "[q1] 2020-01-07 10:00:02,050 [thr][sess:4455667] INFO  CommonTiming::Stop Svc elapsed time = 50 ms"])
    # the exit saves the number from the BAF info; the record was already waiting without one and goes out at expiry
    assert out == ["tx|jvm01|Svc|q1|4455667|1578391202000|1578391202050|50|N"]


def test_ejb_forms_only_in_server_log_and_top_level_flag():
    lines = ["[e1] 2020-01-07 10:00:00,000 INFO  [CommonTiming] The EJB a b c d e f Ledger is starting",
             "[e1] 2020-01-07 10:00:00,250 INFO  [CommonTiming] Total time for the Ledger was 250 ms",
             "[e2] 2020-01-07 10:00:01,000 INFO  [CommonTiming] Total time for the Orphan was 10 ms"]
    assert run(SRV, lines) == ["tx|jvm02|S:Ledger|e1|NaN|1578391200000|1578391200250|250|Y"]  # the orphan exit is dropped (:418)
    assert run(FP, lines) == []  # an application log has no EJB forms (:777-787)


def test_second_start_overwrites_and_second_stop_finds_nothing():
    out = run(FP, ["[k] 2020-01-07 10:00:00,000 INFO  CommonTiming::Start Svc",
                   "[k] 2020-01-07 10:00:01,000 INFO  CommonTiming::Start Svc",
                   "[k] 2020-01-07 10:00:02,000 INFO  CommonTiming::Stop Svc elapsed time = 1000 ms",
                   "[k] 2020-01-07 10:00:03,000 INFO  CommonTiming::Stop Svc elapsed time = 7 ms"])
    assert out == ["tx|jvm01|Svc||NaN|1578391202993|1578391203000|7|N",          # salvaged at once
                   "tx|jvm01|Svc|k|NaN|1578391201000|1578391202000|1000|N"]      # need-number cache, at expiry


def test_info_inside_the_service_cuts_the_split():
    out = run(FP, ["[k] 2020-01-07 10:00:02,000 INFO  CommonTiming::Stop SvcINFOName elapsed time = 5 ms"])
    # line.split(/INFO/)[1] is "  CommonTiming::Stop Svc": service "Svc", elapsed undefined -> NaN
    assert out == ["tx|jvm01|Svc||NaN|NaN|1578391202000|NaN|N"]


def test_synthetic_log_is_parsed_without_surprises():
    text = make_log(seed=3, n=300).decode()
    o = LogParserOracle()
    for l in text.split("\n"):
        o.readLine(SRV, l)
    o.flush()
    assert len(o.out) > 200 and all(l.startswith("tx|jvm02|") for l in o.out)
    assert any("|Y" in l for l in o.out) and any("Provider:" in l for l in o.out)
