#!/usr/bin/env python3
"""Regenerates tests/golden/*.json from the reference's own fixtures.

Run in the build container (needs /root/reference, which does not exist on the GPU box):
    python tests/golden/make_golden.py

Inputs (data fixtures of the reference, copied as data — they carry the reference's own labels):
    pkg/correlation/testdata/labeled_pairs.jsonl       55 labelled span/signal pairs (expected_match, expected_tier)
    pkg/attribution/testdata/multi_fault_samples.jsonl  55 incidents with signals and expected domains
    pkg/benchmark/testdata/mixed_fault_samples.jsonl    4 label-only incidents (rule path)
Outputs:
    labeled_pairs.json, multi_fault_samples.json, mixed_fault_samples.json : the records, verbatim
    scenario_assertions.json : fault_profile and assertions of test/incident-lab/scenarios/*.yaml
    derived.json : results of oracle/ref_str.py on those records, kept as a regression pin next to
                   the independently extracted values of SURVEY.md §4 (checked in tests/test_oracle_str.py)
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("LSG_REFERENCE", "/root/reference")
sys.path.insert(0, ROOT)


def read_jsonl(path):
    with open(path) as f:
        return [json.loads(line) for line in f if line.strip()]


def main():
    pairs = read_jsonl(os.path.join(REF, "pkg/correlation/testdata/labeled_pairs.jsonl"))
    multi = read_jsonl(os.path.join(REF, "pkg/attribution/testdata/multi_fault_samples.jsonl"))
    mixed = read_jsonl(os.path.join(REF, "pkg/benchmark/testdata/mixed_fault_samples.jsonl"))
    for name, data in (("labeled_pairs", pairs), ("multi_fault_samples", multi), ("mixed_fault_samples", mixed)):
        with open(os.path.join(HERE, name + ".json"), "w") as f:
            json.dump(data, f, indent=0, sort_keys=True)
            f.write("\n")

    from tests.fixtures import pairs_from_records, samples_from_records
    from oracle import ref_str as R

    lp = pairs_from_records(pairs)
    report, preds = R.EvaluateLabeledPairs(lp, 2000 * R.MS, 0.7)
    samples = samples_from_records(multi)
    attr = R.BuildAttributionsBayes(samples, None)
    derived = {
        "pairs": {
            "tp": report.TruePositive, "fp": report.FalsePositive, "fn": report.FalseNegative, "tn": report.TrueNegative,
            "precision": report.Precision, "recall": report.Recall, "f1": report.F1,
            "tier_accuracy": report.TierAccuracy, "mean_confidence": report.MeanConfidence,
            "tiers": [p.Tier for p in preds], "predicted": [p.Predicted for p in preds],
        },
        "multi_fault": {
            "predicted": [a.PredictedFaultDomain for a in attr],
            "confidence": [a.Confidence for a in attr],
            "hypotheses": [{h[0]: h[1] for h in a.FaultHypotheses} for a in attr],
            "accuracy": R.Accuracy(samples, attr),
            "partial": R.PartialAccuracy(samples, attr),
            "coverage_0.10": R.CoverageAccuracy(samples, attr, 0.10),
            "confusion": sorted([[k[0], k[1], v] for k, v in R.BuildConfusionMatrix(samples, attr).items()]),
        },
    }
    with open(os.path.join(HERE, "derived.json"), "w") as f:
        json.dump(derived, f, indent=1, sort_keys=True)
        f.write("\n")
    # the incident-lab scenarios' fault profiles and `assertions:` blocks (test/incident-lab/scenarios/*.yaml), as data
    import glob
    import yaml
    scen = {}
    for path in sorted(glob.glob(os.path.join(REF, "test/incident-lab/scenarios/*.yaml"))):
        with open(path) as f:
            doc = yaml.safe_load(f)
        scen[doc["name"]] = {"fault_profile": doc.get("fault_profile", {}), "assertions": doc.get("assertions", []),
                             "sample_count": (doc.get("harness") or {}).get("sample_count")}
    with open(os.path.join(HERE, "scenario_assertions.json"), "w") as f:
        json.dump(scen, f, indent=1, sort_keys=True)
        f.write("\n")
    print("pairs:", {k: derived["pairs"][k] for k in ("tp", "fp", "fn", "tn", "mean_confidence", "tier_accuracy")})
    print("multi:", {k: derived["multi_fault"][k] for k in ("accuracy", "partial", "coverage_0.10")})
    print(derived["multi_fault"]["confusion"])


if __name__ == "__main__":
    main()
