"""Dictionary-encoding of the reference's string-keyed records into the packed 32-byte records of
include/llmslo_gpu.h, including the join-group assignment the C ABI's contract asks for.

This is data preparation (what pkg/gpu does on the Go side), not path arithmetic:
no Match(), no scoring and no attribution happens here.
"""
from datetime import datetime, timezone

import numpy as np

from . import _abi as A


def to_ns(ts):
    """time.Time -> Unix ns (int), None/zero Time -> LSG_TS_ZERO."""
    if ts is None:
        return A.TS_ZERO
    if isinstance(ts, datetime):
        if ts.tzinfo is None:
            ts = ts.replace(tzinfo=timezone.utc)
        if ts.year == 1 and ts.month == 1 and ts.day == 1 and ts.hour == 0 and ts.minute == 0 and ts.second == 0 \
                and ts.microsecond == 0:
            return A.TS_ZERO
        epoch = datetime(1970, 1, 1, tzinfo=timezone.utc)
        d = ts - epoch
        return (d.days * 86400 + d.seconds) * 1_000_000_000 + d.microseconds * 1000
    return int(ts)


class _Dict:
    def __init__(self):
        self.ids = {"": 0}

    def get(self, s):
        i = self.ids.get(s)
        if i is None:
            i = len(self.ids)
            self.ids[s] = i
        return i


class _UnionFind:
    def __init__(self):
        self.p = []

    def add(self):
        self.p.append(len(self.p))
        return len(self.p) - 1

    def find(self, x):
        p = self.p
        while p[x] != x:
            p[x] = p[p[x]]
            x = p[x]
        return x

    def union(self, a, b):
        ra, rb = self.find(a), self.find(b)
        if ra != rb:
            self.p[max(ra, rb)] = min(ra, rb)


class Interner:
    """One interner per batch (spans and signals are interned together so that ids agree)."""

    def __init__(self):
        self.trace, self.pod, self.node, self.service, self.conn = _Dict(), _Dict(), _Dict(), _Dict(), _Dict()
        self.ident_rows = {}       # (pod, node, service) -> row
        self.ident_list = []
        self._uf = _UnionFind()
        self._by_pod, self._by_svcnode, self._by_trace = {}, {}, {}

    def _ident(self, pod, node, service):
        key = (self.pod.get(pod), self.node.get(node), self.service.get(service))
        row = self.ident_rows.get(key)
        if row is None:
            row = len(self.ident_list)
            self.ident_rows[key] = row
            self.ident_list.append(key)
            self._uf.add()
            # rows that could match through tiers 2-3 (same pod) or tier 4 (same service+node): dns.go:62-73
            if key[0] != 0:
                other = self._by_pod.setdefault(key[0], row)
                self._uf.union(other, row)
            if key[1] != 0 and key[2] != 0:
                other = self._by_svcnode.setdefault((key[2], key[1]), row)
                self._uf.union(other, row)
        return row

    def _link_trace(self, trace_id, row):
        # rows that could match through tier 1 (same trace id, any pod): dns.go:58-60
        if trace_id != 0:
            other = self._by_trace.setdefault(trace_id, row)
            self._uf.union(other, row)

    @staticmethod
    def _pid(pid):
        pid = int(pid)
        if pid <= 0:
            return 0
        if pid >= 1 << 32:
            raise ValueError("PID %d does not fit the packed record" % pid)
        return pid

    def span_records(self, spans, incidents=None):
        out = np.zeros(len(spans), A.SPAN_DTYPE)
        for i, s in enumerate(spans):
            row = self._ident(s.Pod, s.Node, s.Service)
            t = self.trace.get(s.TraceID)
            if t >= 1 << A.TRACE_BITS:
                raise ValueError("more than 2^28 distinct trace ids in one batch")
            self._link_trace(t, row)
            out[i] = (to_ns(s.Timestamp), row, self._pid(s.PID), self.conn.get(s.ConnTuple), t,
                      A.NONE if incidents is None else incidents[i], 0)
        return out

    def event_records(self, signals):
        out = np.zeros(len(signals), A.EVENT_DTYPE)
        for i, s in enumerate(signals):
            row = self._ident(s.Pod, s.Node, s.Service)
            t = self.trace.get(s.TraceID)
            if t >= 1 << A.TRACE_BITS:
                raise ValueError("more than 2^28 distinct trace ids in one batch")
            self._link_trace(t, row)
            sig = A.SIGNAL_INDEX.get(s.Signal, A.SIG_UNSUPPORTED)
            out[i] = (to_ns(s.Timestamp), float(s.Value), row, self._pid(s.PID), self.conn.get(s.ConnTuple),
                      (t << 4) | sig)
        return out

    def ident_table(self):
        """Identity rows with their join group: 1 + index of the union-find component."""
        n = len(self.ident_list)
        out = np.zeros(max(n, 1), A.IDENT_DTYPE)
        comp = {}
        for row, (pod, node, service) in enumerate(self.ident_list):
            root = self._uf.find(row)
            g = comp.setdefault(root, len(comp) + 1)
            out[row] = (pod, node, service, g)
        return out
