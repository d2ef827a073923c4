"""Helpers shared by the tests: reference fixture records -> oracle/ref_str types, RFC3339 -> Unix ns."""
import json
import os
import re
from datetime import datetime, timezone

from oracle import ref_str as R

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

_TS = re.compile(r"^(\d{4})-(\d\d)-(\d\d)T(\d\d):(\d\d):(\d\d)(?:\.(\d{1,9}))?Z$")


def parse_ts(s):
    """RFC3339 (UTC, up to ns) -> Unix ns; None for Go's zero Time ("0001-01-01T00:00:00Z") or a missing field."""
    if s is None or s == "" or s.startswith("0001-01-01T00:00:00"):
        return None
    m = _TS.match(s)
    if not m:
        raise ValueError("unsupported timestamp %r" % s)
    y, mo, d, h, mi, sec, frac = m.groups()
    base = int(datetime(int(y), int(mo), int(d), int(h), int(mi), int(sec), tzinfo=timezone.utc).timestamp())
    ns = int((frac or "0").ljust(9, "0"))
    return base * 1_000_000_000 + ns


def span_from(d):
    return R.SpanRef(d.get("trace_id", ""), d.get("service", ""), d.get("node", ""), d.get("pod", ""),
                     int(d.get("pid", 0)), d.get("conn_tuple", ""), parse_ts(d.get("timestamp")))


def signal_from(d):
    return R.SignalRef(d.get("signal", ""), d.get("trace_id", ""), d.get("service", ""), d.get("node", ""),
                       d.get("pod", ""), int(d.get("pid", 0)), d.get("conn_tuple", ""), parse_ts(d.get("timestamp")),
                       float(d.get("value", 0.0)))


def pairs_from_records(records):
    return [R.LabeledPair(r["case_id"], span_from(r["span"]), signal_from(r["signal"]), bool(r["expected_match"]),
                          r.get("expected_tier", "")) for r in records]


def samples_from_records(records):
    return [R.FaultSample(r.get("incident_id", ""), r.get("fault_label", ""), r.get("expected_domain", ""),
                          list(r.get("expected_domains") or []), dict(r.get("signals") or {}),
                          float(r.get("confidence", 0.0))) for r in records]


def load_golden(name):
    with open(os.path.join(GOLDEN, name + ".json")) as f:
        return json.load(f)
