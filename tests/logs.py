"""Synthetic WildFly log text for the log-feeder tests (formats per stream_parse_transactions.js:378-565)."""
import random

SERVICES = ["AccountLookup", "Provider[acct].get", "provider[Risk]Score", "S:TopLevel", "Quote]Odd", "BalanceService",
            "Transfer.execute", "a", "Login"]


def stamp(ms):
    """epoch ms -> `YYYY-MM-DD HH:MM:SS,mmm` in UTC (the tests run the parser with a zero offset)."""
    import datetime
    d = datetime.datetime.fromtimestamp(ms / 1000.0, datetime.timezone.utc)
    return d.strftime("%Y-%m-%d %H:%M:%S,") + f"{ms % 1000:03d}"


def make_log(seed=1, n=400, t0=1_578_391_200_000, server_log=True, span_ms=120_000, unique_ids=False):
    """Interleaved Start / Stop (and, for server.log, EJB entry / exit) lines with the oddities the parser
    has to survive: empty log ids, BAF meta info before INFO, unknown lines, stops without a start, starts
    without a stop, repeated (logId, service), a service name containing INFO, CRLF, no final newline."""
    rnd = random.Random(seed)
    events = []
    for i in range(n):
        svc = rnd.choice(SERVICES)
        log_id = "" if rnd.random() < 0.05 else (f"u{i:06d}" if unique_ids else f"id{rnd.randrange(n // 2):05d}")
        start = t0 + rnd.randrange(span_ms)
        elapsed = rnd.choice([0, 1, 7, 35, 120, 999, 4321, 65536, 70000])
        ejb = server_log and rnd.random() < 0.3
        baf = "" if rnd.random() < 0.8 else f" [x][sess:{rnd.randrange(10**6)}]" if rnd.random() < 0.5 else " [sess:abc]"
        has_start, has_stop = rnd.random() > 0.06, rnd.random() > 0.06
        if ejb:
            if has_start:
                events.append((start, f"[{log_id}] {stamp(start)} INFO  [CommonTiming] The EJB a b c d e f {svc} is starting"))
            if has_stop:
                events.append((start + elapsed, f"[{log_id}] {stamp(start + elapsed)} INFO  [CommonTiming] Total time for the {svc} was {elapsed} ms"))
        else:
            thread = f"[thr-{rnd.randrange(8)}]"
            if has_start:
                events.append((start, f"[{log_id}] {stamp(start)} {thread}{baf} INFO  CommonTiming::Start {svc}"))
            if has_stop:
                events.append((start + elapsed, f"[{log_id}] {stamp(start + elapsed)} {thread}{baf} INFO  CommonTiming::Stop {svc} elapsed time = {elapsed} ms"))
    for i in range(n // 10):
        t = t0 + rnd.randrange(span_ms)
        events.append((t, rnd.choice([
            f"[idz] {stamp(t)} [thr-1] DEBUG something else entirely",
            f"{stamp(t)} INFO no log id here",
            "",
            "   ",
            f"[idq] {stamp(t)} [thr-2] INFO  CommonTiming::Stop SvcINFOName elapsed time = 5 ms",
            f"[idq] {stamp(t)} [thr-2] INFO  CommonTiming::Stop Login elapsed time = soon ms",
            f"[idq] not-a-date nor-time [thr-2] INFO  CommonTiming::Stop Login elapsed time = 12 ms",
            f"[idw] {stamp(t)} [thr-2] INFO CommonTiming::Stop",
            f"  [idv] {stamp(t)} [thr-2] INFO  CommonTiming::Start Login",
            f"[idu] {stamp(t)} [thr-2] INFO  CommonTiming::Stop Login elapsed time = 44 ms\r",
        ])))
    events.sort(key=lambda e: e[0])
    text = "\n".join(line for _, line in events)
    return text.encode()
