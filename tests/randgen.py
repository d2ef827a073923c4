"""Randomised string-keyed spans/signals that exercise every tier, both window edges, ties, empty
fields, zero times, unsupported signal types and cross-pod traces."""
import random

from oracle import ref_str as R

MS, S = R.MS, R.SECOND
BASE = 1_771_372_800_000_000_000
EDGES_MS = [0, 1, 50, 99, 100, 101, 249, 250, 251, 499, 500, 501, 1999, 2000, 2001, 3100]


def _pick(rng, items, p_empty=0.15):
    return "" if rng.random() < p_empty else rng.choice(items)


def random_batch(seed, n_spans, n_signals, n_pods=3, window_ms=2000, dense=True):
    rng = random.Random(seed)
    pods = ["pod-%d" % i for i in range(n_pods)]
    nodes = ["node-%d" % i for i in range(max(1, n_pods // 2))]
    svcs = ["svc-%d" % i for i in range(2)]
    traces = ["trace-%d" % i for i in range(max(2, n_spans // 2))]
    conns = ["10.0.0.%d:4%03d->10.0.0.53:443/tcp" % (i, i) for i in range(3)]
    pids = [0, 0, 101, 102, 103, -5]
    horizon = (3 if dense else 30) * S
    span_ts = []
    spans = []
    for _ in range(n_spans):
        pod = _pick(rng, pods)
        # most pods have a canonical node/service, some records disagree (fixture c49)
        node = _pick(rng, nodes, 0.1) if rng.random() < 0.2 else (nodes[pods.index(pod) % len(nodes)] if pod else _pick(rng, nodes))
        svc = _pick(rng, svcs, 0.1)
        ts = None if rng.random() < 0.03 else BASE + rng.randrange(0, horizon // MS) * MS
        span_ts.append(ts if ts is not None else BASE)
        spans.append(R.SpanRecord(_pick(rng, traces, 0.2), "", svc, node, pod, rng.choice(pids), _pick(rng, conns, 0.3), ts,
                                  None if rng.random() < 0.8 else {R.ATTR["dns_latency_ms"]: rng.choice([5.0, 500.0]),
                                                                  "custom.attr": 1.0}))
    signals = []
    names = R.SIGNAL_ORDER + ["unknown_signal"]
    for _ in range(n_signals):
        anchor = rng.randrange(n_spans) if n_spans else 0
        sp = spans[anchor] if n_spans else R.SpanRecord()
        mode = rng.random()
        if mode < 0.6 and n_spans:   # derived from a span: shares some identity, time near an edge
            dt = rng.choice(EDGES_MS) * MS * rng.choice([-1, 1])
            if rng.random() < 0.3:
                dt += rng.choice([-1, 0, 1])           # +-1 ns around the edge
            ts = span_ts[anchor] + dt
            sig = R.SignalRef(rng.choice(names),
                              sp.TraceID if rng.random() < 0.5 else _pick(rng, traces, 0.5),
                              sp.Service if rng.random() < 0.7 else _pick(rng, svcs),
                              sp.Node if rng.random() < 0.7 else _pick(rng, nodes),
                              sp.Pod if rng.random() < 0.7 else _pick(rng, pods),
                              sp.PID if rng.random() < 0.6 else rng.choice(pids),
                              sp.ConnTuple if rng.random() < 0.6 else _pick(rng, conns, 0.3),
                              ts, rng.choice([0.0, 1.5, 12.0, 99.0, 250.0, -3.0]))
        else:
            ts = None if rng.random() < 0.03 else BASE + rng.randrange(0, horizon // MS) * MS
            sig = R.SignalRef(rng.choice(names), _pick(rng, traces, 0.4), _pick(rng, svcs), _pick(rng, nodes),
                              _pick(rng, pods), rng.choice(pids), _pick(rng, conns, 0.3), ts,
                              rng.choice([0.0, 1.5, 12.0, 99.0, 250.0]))
        signals.append(sig)
    return spans, signals
