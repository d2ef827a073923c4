//go:build gpu

package correlation

import (
	"time"

	"github.com/ogulcanaydogan/llm-slo-ebpf-toolkit/pkg/gpu"
)

func tierCode(name string) uint8 {
	for i, t := range gpu.Tiers {
		if i > 0 && t == name {
			return uint8(i)
		}
	}
	return 0
}

// EvaluateLabeledPairs computes quality metrics at a given threshold/window (GPU build: evaluator.go:90-167 with
// the Match loop replaced by one lsg_match_pairs call; the report's ratios are formed here from integer counts).
func EvaluateLabeledPairs(pairs []LabeledPair, window time.Duration, threshold float64) (EvalReport, []Prediction) {
	if window <= 0 {
		window = DefaultWindow
	}
	if threshold <= 0 {
		threshold = DefaultEnrichmentThreshold
	}
	report := EvalReport{GeneratedAt: time.Now().UTC(), SampleSize: len(pairs), WindowMS: int(window / time.Millisecond),
		Threshold: threshold}
	predictions := make([]Prediction, 0, len(pairs))
	if len(pairs) == 0 {
		return report, predictions
	}

	in := gpu.NewInterner()
	spans := make([]gpu.Span, len(pairs))
	events := make([]gpu.Event, len(pairs))
	expMatch := make([]uint8, len(pairs))
	expTier := make([]uint8, len(pairs))
	for i, p := range pairs {
		var err error
		if spans[i], err = in.Span(p.Span.TraceID, p.Span.Service, p.Span.Node, p.Span.Pod, p.Span.PID, p.Span.ConnTuple,
			p.Span.Timestamp, gpu.None); err != nil {
			panic(err)
		}
		if events[i], err = in.Event(p.Signal.Signal, p.Signal.TraceID, p.Signal.Service, p.Signal.Node, p.Signal.Pod,
			p.Signal.PID, p.Signal.ConnTuple, p.Signal.Timestamp, p.Signal.Value); err != nil {
			panic(err)
		}
		if p.ExpectedMatch {
			expMatch[i] = 1
		}
		// an expected tier that is not one of Match's four names can never equal decision.Tier: code 7 keeps the
		// pair comparable (evaluator.go:143-148) and never correct
		if p.ExpectedTier != "" {
			if expTier[i] = tierCode(p.ExpectedTier); expTier[i] == 0 {
				expTier[i] = 7
			}
		}
	}
	tier, predicted, ev, err := gpu.MustDefault().MatchPairs(spans, events, in.Idents(), int64(window), threshold, expMatch, expTier)
	if err != nil {
		panic(err) // a broken device; the reference's signature has no error to return
	}

	for i, p := range pairs {
		pred := predicted[i] != 0
		predictions = append(predictions, Prediction{CaseID: p.CaseID, Expected: p.ExpectedMatch, Predicted: pred,
			Confidence: gpu.TierConfidence[tier[i]], Tier: gpu.Tiers[tier[i]], Correct: pred == p.ExpectedMatch,
			Signal: p.Signal.Signal, ExpectedTier: p.ExpectedTier})
	}
	report.TruePositive, report.FalsePositive = int(ev.TP), int(ev.FP)
	report.FalseNegative, report.TrueNegative = int(ev.FN), int(ev.TN)
	report.Precision = safeDiv(report.TruePositive, report.TruePositive+report.FalsePositive)
	report.Recall = safeDiv(report.TruePositive, report.TruePositive+report.FalseNegative)
	if report.Precision+report.Recall > 0 {
		report.F1 = 2 * ((report.Precision * report.Recall) / (report.Precision + report.Recall))
	}
	if ev.TierComparable > 0 {
		report.TierAccuracy = float64(ev.TierCorrect) / float64(ev.TierComparable)
	}
	// the reference adds the predicted pairs' confidences in input order; the confidences are four literals, so
	// the same sum is formed here in input order from the per-pair tiers
	confSum, confCount := 0.0, 0
	for i := range pairs {
		if predicted[i] != 0 {
			confSum += gpu.TierConfidence[tier[i]]
			confCount++
		}
	}
	if confCount > 0 {
		report.MeanConfidence = confSum / float64(confCount)
	}
	return report, predictions
}
