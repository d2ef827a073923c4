"""Artifact writers byte-compatible with the reference CLIs, so that the files the CI gates
diff — predictions.csv / eval_summary.json of cmd/correlationeval and confusion.csv / summary.json of
cmd/attributor — come out the same when the numbers come from the GPU path.  Pure host-side formatting."""
import csv
import io
import json
import math
import os
from datetime import datetime, timezone


def go_float(f: float) -> str:
    """encoding/json's float64 encoding: the shortest digits that round-trip, positional ('f') unless the exponent is
    < -6 or >= 21, then 'e' with Go's exponent clean-up (e-09 -> e-9)."""
    from decimal import Decimal
    f = float(f)
    if math.isnan(f) or math.isinf(f):
        raise ValueError("json: unsupported value: %r" % f)
    if f == 0:
        return "-0" if math.copysign(1.0, f) < 0 else "0"
    sign, digits, exp = Decimal(repr(f)).as_tuple()
    digits = "".join(map(str, digits)).lstrip("0")
    while len(digits) > 1 and digits.endswith("0"):
        digits, exp = digits[:-1], exp + 1
    nd = len(digits)
    dp = nd + exp                       # value = 0.DIGITS x 10^dp
    a = abs(f)
    if a < 1e-6 or a >= 1e21:
        e10 = dp - 1
        mant = digits[0] + ("." + digits[1:] if nd > 1 else "")
        es = "%02d" % abs(e10)
        if e10 < 0 and es[0] == "0":
            es = es[1:]                 # encoding/json: "clean up e-09 to e-9"
        out = "%se%s%s" % (mant, "-" if e10 < 0 else "+", es)
    elif dp <= 0:
        out = "0." + "0" * (-dp) + digits
    elif dp >= nd:
        out = digits + "0" * (dp - nd)
    else:
        out = digits[:dp] + "." + digits[dp:]
    return ("-" if sign else "") + out


def _go_json(obj, indent="  ", level=0) -> str:
    """json.MarshalIndent for dicts (insertion order = struct field order), lists, str, bool, int, float."""
    pad, pad_in = indent * level, indent * (level + 1)
    if isinstance(obj, dict):
        if not obj:
            return "{}"
        items = ["%s%s: %s" % (pad_in, json.dumps(k), _go_json(v, indent, level + 1)) for k, v in obj.items()]
        return "{\n" + ",\n".join(items) + "\n" + pad + "}"
    if isinstance(obj, (list, tuple)):
        if not obj:
            return "[]"
        return "[\n" + ",\n".join(pad_in + _go_json(v, indent, level + 1) for v in obj) + "\n" + pad + "]"
    if isinstance(obj, bool):
        return "true" if obj else "false"
    if isinstance(obj, int):
        return str(obj)
    if isinstance(obj, float):
        return go_float(obj)
    if obj is None:
        return "null"
    return json.dumps(obj, ensure_ascii=False)


def _bool(v) -> str:   # strconv.FormatBool
    return "true" if v else "false"


def _csv(rows) -> str:
    buf = io.StringIO()
    w = csv.writer(buf, lineterminator="\n", quoting=csv.QUOTE_MINIMAL)   # encoding/csv: quote only when needed, "\n"
    for r in rows:
        w.writerow(r)
    return buf.getvalue()


def _write(path, text):
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    with open(path, "w", newline="") as f:
        f.write(text)


def PredictionsCSV(predictions) -> str:
    """cmd/correlationeval/main.go:97-136"""
    rows = [["case_id", "signal", "expected_match", "predicted_match", "confidence", "tier", "expected_tier", "is_correct"]]
    for p in predictions:
        rows.append([p.CaseID, p.Signal, _bool(p.Expected), _bool(p.Predicted), "%.4f" % p.Confidence, p.Tier, p.ExpectedTier,
                     _bool(p.Correct)])
    return _csv(rows)


def EvalSummaryJSON(report, generated_at=None) -> str:
    """cmd/correlationeval/main.go:86-95: json.MarshalIndent(EvalReport) with the struct's field order and omitempty rules
    (pkg/correlation/evaluator.go:34-51)."""
    ts = generated_at or datetime.now(timezone.utc).strftime("%Y-%m-%dT%H:%M:%S.%fZ")
    d = {"generated_at": ts, "sample_size": report.SampleSize, "true_positive": report.TruePositive,
         "false_positive": report.FalsePositive, "false_negative": report.FalseNegative, "true_negative": report.TrueNegative,
         "precision": float(report.Precision), "recall": float(report.Recall), "f1": float(report.F1),
         "tier_accuracy": float(report.TierAccuracy), "mean_confidence": float(report.MeanConfidence),
         "window_ms": report.WindowMS, "threshold": float(report.Threshold)}
    if getattr(report, "MinPrecisionReq", 0):
        d["min_precision_required"] = float(report.MinPrecisionReq)
    if getattr(report, "MinRecallReq", 0):
        d["min_recall_required"] = float(report.MinRecallReq)
    if getattr(report, "PassedGate", False):
        d["passed_gate"] = True
    return _go_json(d)


def ConfusionCSV(matrix) -> str:
    """cmd/attributor/main.go:243-283: rows sorted by (actual, predicted).  `matrix` maps MatrixKey -> count."""
    keys = sorted(matrix, key=lambda k: (k.Actual, k.Predicted))
    return _csv([["actual", "predicted", "count"]] + [[k.Actual, k.Predicted, "%d" % matrix[k]] for k in keys])


def AttributorSummaryJSON(samples, predictions, accuracy, attribution_mode="bayes", input_path="", output_path="",
                          confusion_path="", webhook_enabled=False, webhook_strict=False, webhook_errors=0, generated_at=None) -> str:
    """cmd/attributor/main.go:284-327 (summaryPayload); map keys are sorted as encoding/json does."""
    counts = {}
    for p in predictions:
        counts[p.PredictedFaultDomain] = counts.get(p.PredictedFaultDomain, 0) + 1
    d = {"generated_at": generated_at or datetime.now(timezone.utc).strftime("%Y-%m-%dT%H:%M:%SZ"),
         "total_samples": len(samples), "accuracy": float(accuracy), "attribution_mode": attribution_mode,
         "predicted_domain_counts": dict(sorted(counts.items())), "webhook_enabled": bool(webhook_enabled),
         "webhook_strict": bool(webhook_strict), "webhook_delivery_errors": int(webhook_errors)}
    for key, v in (("input_path", input_path), ("output_path", output_path), ("confusion_path", confusion_path)):
        if v:                           # `json:",omitempty"`
            d[key] = v
    return _go_json(d)


def WritePredictionsCSV(path, predictions):
    _write(path, PredictionsCSV(predictions))


def WriteEvalSummary(path, report, generated_at=None):
    _write(path, EvalSummaryJSON(report, generated_at))


def WriteConfusionCSV(path, matrix):
    _write(path, ConfusionCSV(matrix))


# ---- cmd/attributor's attributions.jsonl and pkg/benchmark's harness artifacts ---------------------------------------
def _go_json_compact(obj) -> str:
    """json.Marshal / json.Encoder.Encode without indentation: struct field order, Go float formatting, HTML-safe
    escaping of <, > and & as encoding/json does by default."""
    if isinstance(obj, dict):
        return "{" + ",".join("%s:%s" % (_go_json_compact(str(k)), _go_json_compact(v)) for k, v in obj.items()) + "}"
    if isinstance(obj, (list, tuple)):
        return "[" + ",".join(_go_json_compact(v) for v in obj) + "]"
    if isinstance(obj, bool):
        return "true" if obj else "false"
    if isinstance(obj, int):
        return str(obj)
    if isinstance(obj, float):
        return go_float(obj)
    if obj is None:
        return "null"
    return json.dumps(obj, ensure_ascii=False).replace("<", "\\u003c").replace(">", "\\u003e").replace("&", "\\u0026")


def _rfc3339(ts) -> str:
    """time.Time.Format(time.RFC3339) of a sample timestamp given as an RFC3339 string (fractions are dropped, as
    RFC3339 without Nano does) or None / empty for Go's zero Time."""
    if not ts:
        return "0001-01-01T00:00:00Z"
    return ts.split(".")[0] + ("Z" if "." in ts and not ts.split(".")[0].endswith("Z") else "")


def _rfc3339_minus(ts, seconds) -> str:
    if not ts:
        # zero Time minus 30 s: year 0, which Go prints as 0000-12-31T23:59:30Z
        return "0000-12-31T23:59:%02dZ" % (60 - seconds)
    from datetime import timedelta
    base = datetime.strptime(_rfc3339(ts), "%Y-%m-%dT%H:%M:%SZ").replace(tzinfo=timezone.utc)
    return (base - timedelta(seconds=seconds)).strftime("%Y-%m-%dT%H:%M:%SZ")


def _json_time(ts) -> str:
    """time.Time's MarshalJSON (RFC3339Nano: trailing zeros of the fraction dropped)."""
    if not ts:
        return "0001-01-01T00:00:00Z"
    if "." in ts:
        head, frac = ts.rstrip("Z").split(".")
        frac = frac.rstrip("0")
        return head + ("." + frac if frac else "") + "Z"
    return ts


def AttributionsJSONL(predictions) -> str:
    """cmd/attributor/main.go:203-224: one json.Encoder.Encode(schema.IncidentAttribution) per line
    (pkg/schema/types.go:20-57 field order and omitempty rules)."""
    lines = []
    for p in predictions:
        d = {"incident_id": p.IncidentID, "timestamp": _json_time(p.Timestamp), "cluster": p.Cluster}
        if p.Namespace:
            d["namespace"] = p.Namespace
        d["service"] = p.Service
        d["predicted_fault_domain"] = p.PredictedFaultDomain
        d["confidence"] = float(p.Confidence)
        d["evidence"] = [{"signal": e["signal"], "value": float(e["value"]) if isinstance(e["value"], (int, float)) and
                          not isinstance(e["value"], bool) else e["value"], "source": e["source"]} for e in p.Evidence]
        d["slo_impact"] = {"sli": p.SLOImpact.get("sli", ""), "burn_rate": float(p.SLOImpact.get("burn_rate", 0.0)),
                           "window_minutes": int(p.SLOImpact.get("window_minutes", 0))}
        if p.TraceIDs:
            d["trace_ids"] = list(p.TraceIDs)
        if p.RequestIDs:
            d["request_ids"] = list(p.RequestIDs)
        if p.FaultHypotheses:
            d["fault_hypotheses"] = [{"domain": h.Domain, "posterior": float(h.Posterior), "evidence": list(h.Evidence)}
                                     for h in p.FaultHypotheses]
        lines.append(_go_json_compact(d) + "\n")
    return "".join(lines)


def IncidentPredictionsCSV(scenario, samples, predictions, map_fault_label) -> str:
    """pkg/benchmark/harness.go:285-343 (writeIncidentPredictions): "%.2f" confidence, the constant 2.5 s delay."""
    rows = [["timestamp", "incident_id", "scenario", "fault_start_ts", "fault_end_ts", "predicted_fault_domain",
             "ground_truth_fault_domain", "confidence", "detection_delay_seconds", "is_correct"]]
    for sample, pred in zip(samples, predictions):
        actual = sample.ExpectedDomain or map_fault_label(sample.FaultLabel)
        rows.append([_rfc3339(sample.Timestamp), sample.IncidentID, scenario, _rfc3339_minus(sample.Timestamp, 30),
                     _rfc3339(sample.Timestamp), pred.PredictedFaultDomain, actual, "%.2f" % pred.Confidence, "2.5",
                     _bool(pred.PredictedFaultDomain == actual)])
    return _csv(rows)


def HarnessConfusionCSV(matrix) -> str:
    """pkg/benchmark/harness.go:345-377 (writeConfusionMatrix): the same rows as cmd/attributor's confusion.csv."""
    return ConfusionCSV(matrix)


def WriteAttributionsJSONL(path, predictions):
    _write(path, AttributionsJSONL(predictions))


def WriteIncidentPredictions(path, scenario, samples, predictions, map_fault_label):
    _write(path, IncidentPredictionsCSV(scenario, samples, predictions, map_fault_label))


def WriteAttributorSummary(path, *args, **kw):
    _write(path, AttributorSummaryJSON(*args, **kw))
