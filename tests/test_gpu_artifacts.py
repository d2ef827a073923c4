"""F3 on the GPU: the reference CLIs' artifacts written from GPU results for both reference fixtures, against the
same writers fed by the string oracle (oracle/ref_str.py).

cmd/correlationeval: predictions.csv, eval_summary.json (main.go:86-136); cmd/attributor: attributions.jsonl,
confusion.csv, summary.json (main.go:203-327); pkg/benchmark: incident predictions and confusion matrix
(harness.go:285-377).  Byte for byte wherever the reference's formatting fixes the digits ("%.4f", "%.2f", integers,
labels); fp64 fields that are only 1e-6-comparable by the parity contract (mean confidence, posteriors) are
compared as numbers and everything around them as bytes."""
import json
import os
import re

import pytest

from llm_slo_ebpf_toolkit_b200 import artifacts as AR
from llm_slo_ebpf_toolkit_b200 import attribution as AT
from llm_slo_ebpf_toolkit_b200 import correlation as CO
from oracle import ref_str as R
from tests.fixtures import load_golden, pairs_from_records, parse_ts, samples_from_records

pytestmark = pytest.mark.gpu
MS = 1_000_000


def _gpu_pairs():
    out = []
    for r in load_golden("labeled_pairs"):
        sp, sg = r["span"], r["signal"]
        out.append(CO.LabeledPair(
            r["case_id"],
            CO.SpanRef(sp.get("trace_id", ""), sp.get("service", ""), sp.get("node", ""), sp.get("pod", ""),
                       sp.get("pid", 0), sp.get("conn_tuple", ""), parse_ts(sp.get("timestamp"))),
            CO.SignalRef(sg["signal"], sg.get("trace_id", ""), sg.get("service", ""), sg.get("node", ""), sg.get("pod", ""),
                         sg.get("pid", 0), sg.get("conn_tuple", ""), parse_ts(sg.get("timestamp")), sg.get("value", 0.0)),
            r["expected_match"], r.get("expected_tier", "")))
    return out


def _gpu_samples(name):
    return [AT.FaultSample(r.get("incident_id", ""), r.get("timestamp"), r.get("cluster", ""), r.get("namespace", ""),
                           r.get("service", ""), r.get("fault_label", ""), r.get("expected_domain", ""),
                           list(r.get("expected_domains") or []), dict(r.get("signals") or {}), r.get("confidence", 0.0),
                           r.get("burn_rate", 0.0), r.get("window_minutes", 0), r.get("request_id", ""), r.get("trace_id", ""))
            for r in load_golden(name)]


def _split_number(text, key):
    """(text with the value of `"key": <number>` blanked, the number)"""
    m = re.search(r'("%s": )([-0-9.e+]+)' % re.escape(key), text)
    return text[:m.start(2)] + "#" + text[m.end(2):], float(m.group(2))


def test_correlationeval_artifacts_from_gpu_results(ctx, tmp_path):
    report, preds = CO.EvaluateLabeledPairs(_gpu_pairs(), 2000 * MS, 0.7, ctx)
    want_report, want_preds = R.EvaluateLabeledPairs(pairs_from_records(load_golden("labeled_pairs")), 2000 * MS, 0.7)
    AR.WritePredictionsCSV(str(tmp_path / "gpu" / "predictions.csv"), preds)
    AR.WritePredictionsCSV(str(tmp_path / "ref" / "predictions.csv"), want_preds)
    got = (tmp_path / "gpu" / "predictions.csv").read_bytes()
    assert got == (tmp_path / "ref" / "predictions.csv").read_bytes() and got.count(b"\n") == 56
    stamp = "2026-02-18T00:00:00.000000001Z"
    AR.WriteEvalSummary(str(tmp_path / "gpu" / "eval_summary.json"), report, stamp)
    AR.WriteEvalSummary(str(tmp_path / "ref" / "eval_summary.json"), want_report, stamp)
    g, gm = _split_number((tmp_path / "gpu" / "eval_summary.json").read_text(), "mean_confidence")
    w, wm = _split_number((tmp_path / "ref" / "eval_summary.json").read_text(), "mean_confidence")
    assert g == w and abs(gm - wm) <= 1e-6 * wm
    assert json.loads((tmp_path / "gpu" / "eval_summary.json").read_text())["true_positive"] == 40


def _oracle_envelopes(gpu_samples, ref_attr):
    """The oracle's numbers inside the host-built envelope (BuildAttribution is string work that stays on the host)."""
    out = []
    for s, a in zip(gpu_samples, ref_attr):
        env = AT.BuildAttribution(s)
        env.PredictedFaultDomain, env.Confidence = a.PredictedFaultDomain, a.Confidence
        env.FaultHypotheses = [AT.FaultHypothesis(d, p, list(ev)) for d, p, ev in a.FaultHypotheses]
        out.append(env)
    return out


def test_attributor_and_harness_artifacts_from_gpu_results(ctx, tmp_path):
    for fixture, mode in (("multi_fault_samples", "bayes"), ("mixed_fault_samples", "bayes"), ("multi_fault_samples", "rule")):
        samples = _gpu_samples(fixture)
        attr = AT.BuildAttributions(samples, mode)
        rs = samples_from_records(load_golden(fixture))
        ref_attr = _oracle_envelopes(samples, R.BuildAttributions(rs, mode))
        gdir, rdir = tmp_path / ("gpu_%s_%s" % (fixture, mode)), tmp_path / ("ref_%s_%s" % (fixture, mode))
        # confusion.csv (cmd/attributor) = confusion matrix (pkg/benchmark): counts, byte for byte
        gm = AT.BuildConfusionMatrix(samples, attr, ctx)
        rm = {AT.MatrixKey(a, p): c for (a, p), c in R.BuildConfusionMatrix(rs, R.BuildAttributions(rs, mode)).items()}
        AR.WriteConfusionCSV(str(gdir / "confusion.csv"), gm)
        AR.WriteConfusionCSV(str(rdir / "confusion.csv"), rm)
        assert (gdir / "confusion.csv").read_bytes() == (rdir / "confusion.csv").read_bytes()
        assert AR.HarnessConfusionCSV(gm) == (rdir / "confusion.csv").read_text()
        # summary.json: accuracy is a ratio of two integers on both sides, so the whole file is byte-comparable
        stamp = "2026-02-20T00:00:00Z"
        AR.WriteAttributorSummary(str(gdir / "summary.json"), samples, attr, AT.Accuracy(samples, attr, ctx), mode, "in.jsonl",
                                  "out.jsonl", "confusion.csv", generated_at=stamp)
        AR.WriteAttributorSummary(str(rdir / "summary.json"), samples, ref_attr, R.Accuracy(rs, R.BuildAttributions(rs, mode)), mode,
                                  "in.jsonl", "out.jsonl", "confusion.csv", generated_at=stamp)
        assert (gdir / "summary.json").read_bytes() == (rdir / "summary.json").read_bytes()
        # incident predictions of the harness: "%.2f" confidences
        AR.WriteIncidentPredictions(str(gdir / "incident_predictions.csv"), "mixed_multi", samples, attr, AT.MapFaultLabel)
        AR.WriteIncidentPredictions(str(rdir / "incident_predictions.csv"), "mixed_multi", samples, ref_attr, AT.MapFaultLabel)
        assert (gdir / "incident_predictions.csv").read_bytes() == (rdir / "incident_predictions.csv").read_bytes()
        # attributions.jsonl: every field byte for byte except the fp64 confidence / posteriors (1e-6 relative)
        AR.WriteAttributionsJSONL(str(gdir / "attributions.jsonl"), attr)
        AR.WriteAttributionsJSONL(str(rdir / "attributions.jsonl"), ref_attr)
        glines = (gdir / "attributions.jsonl").read_text().splitlines()
        rlines = (rdir / "attributions.jsonl").read_text().splitlines()
        assert len(glines) == len(rlines) == len(samples)
        num = re.compile(r'"(confidence|posterior)":([-0-9.e+]+)')
        for gl, rl in zip(glines, rlines):
            assert num.sub(r'"\1":#', gl) == num.sub(r'"\1":#', rl)
            for (_, a), (_, b) in zip(num.findall(gl), num.findall(rl)):
                assert abs(float(a) - float(b)) <= 1e-6 * max(abs(float(b)), 1e-300)
            json.loads(gl)
    assert os.path.exists(str(gdir / "attributions.jsonl"))
