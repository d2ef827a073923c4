"""F3: the artifact writers reproduce the reference CLIs' formatting (cmd/correlationeval/main.go:86-136,
cmd/attributor/main.go:243-327): encoding/csv quoting, "%.4f" confidences, strconv.FormatBool, encoding/json float
and indent rules, rows sorted by (actual, predicted).  CPU only: the numbers come from the string oracle here."""
import json

from llm_slo_ebpf_toolkit_b200 import artifacts as AR
from llm_slo_ebpf_toolkit_b200.attribution import MatrixKey
from oracle import ref_str as R
from tests.fixtures import load_golden, pairs_from_records, samples_from_records


def test_go_float_matches_encoding_json():
    cases = {1.0: "1", 0.5: "0.5", 0.92: "0.92", 0.9199999999999997: "0.9199999999999997", 29 / 55: "0.5272727272727272",
             100.0: "100", 1e20: "100000000000000000000", 1e21: "1e+21", 1.5e-7: "1.5e-7", 0.000001: "0.000001", 9.999e-7: "9.999e-7",
             1e-10: "1e-10", 123456789.125: "123456789.125", -2.5: "-2.5", 0.0: "0", 2000.0: "2000", 0.7: "0.7", 3.0e-5: "0.00003"}
    for v, want in cases.items():
        assert AR.go_float(v) == want, (v, AR.go_float(v), want)
        assert float(AR.go_float(v)) == v


def test_predictions_csv_and_summary():
    pairs = pairs_from_records(load_golden("labeled_pairs"))
    report, preds = R.EvaluateLabeledPairs(pairs, 2000 * R.MS, 0.7)
    text = AR.PredictionsCSV(preds)
    lines = text.split("\n")
    assert lines[0] == "case_id,signal,expected_match,predicted_match,confidence,tier,expected_tier,is_correct"
    assert lines[1] == "c01,dns_latency_ms,true,true,1.0000,trace_id_exact,trace_id_exact,true"
    assert len(lines) == 57 and lines[-1] == ""
    none = [ln for ln in lines[1:-1] if ",false,false,0.0000,," in ln]
    assert len(none) == 6                                          # the six unmatched pairs: empty tier and expected tier
    low = [ln for ln in lines[1:-1] if ",0.6500,service_node_500ms," in ln]
    assert len(low) == 9 and all(",false,false,0.6500," in ln for ln in low)
    js = AR.EvalSummaryJSON(report, generated_at="2026-02-18T00:00:00.000000001Z")
    assert js.startswith('{\n  "generated_at": "2026-02-18T00:00:00.000000001Z",\n  "sample_size": 55,\n  "true_positive": 40,')
    assert '"precision": 1,\n  "recall": 1,\n  "f1": 1,\n  "tier_accuracy": 1,\n  "mean_confidence": 0.9199999999999997,' in js
    assert js.endswith('"window_ms": 2000,\n  "threshold": 0.7\n}')     # the three omitempty fields are absent
    assert json.loads(js)["true_negative"] == 15


def test_confusion_csv_and_attributor_summary():
    samples = samples_from_records(load_golden("multi_fault_samples"))
    attr = R.BuildAttributionsBayes(samples, None)
    matrix = {MatrixKey(a, p): c for (a, p), c in R.BuildConfusionMatrix(samples, attr).items()}
    text = AR.ConfusionCSV(matrix)
    assert text.split("\n")[:4] == ["actual,predicted,count", "cpu_throttle,cpu_throttle,6", "memory_pressure,memory_pressure,6",
                                    "network_dns,network_dns,7"]
    assert text.split("\n")[-2] == "unknown,provider_throttle,3"
    weird = AR.ConfusionCSV({MatrixKey('a,b', 'x"y'): 1})
    assert weird == 'actual,predicted,count\n"a,b","x""y",1\n'         # encoding/csv quoting
    js = AR.AttributorSummaryJSON(samples, attr, R.Accuracy(samples, attr), input_path="in.jsonl", generated_at="2026-02-20T00:00:00Z")
    d = json.loads(js)
    assert list(d) == ["generated_at", "total_samples", "accuracy", "attribution_mode", "predicted_domain_counts", "webhook_enabled",
                       "webhook_strict", "webhook_delivery_errors", "input_path"]
    assert '"accuracy": 0.5272727272727272,' in js and list(d["predicted_domain_counts"]) == sorted(d["predicted_domain_counts"])
    assert sum(d["predicted_domain_counts"].values()) == 55
