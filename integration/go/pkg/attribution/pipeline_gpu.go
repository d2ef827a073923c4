//go:build gpu

package attribution

import (
	"math"
	"sort"

	"github.com/ogulcanaydogan/llm-slo-ebpf-toolkit/pkg/gpu"
	"github.com/ogulcanaydogan/llm-slo-ebpf-toolkit/pkg/schema"
)

func domainIndex(name string) int {
	for i, d := range gpu.Domains {
		if d == name {
			return i
		}
	}
	return -1
}

// model flattens b.Priors / b.Likelihoods (public, mutable maps: flattened on every call so edits take effect).
func (b *BayesianAttributor) model() gpu.Model {
	var m gpu.Model
	for d, name := range gpu.Domains {
		m.Prior[d] = b.Priors[name] // missing -> 0 -> 1e-10 inside the library (bayesian.go:231-233)
	}
	for s, name := range gpu.Signals {
		m.Thresh[s] = signalThresholds[name]
		ll, ok := b.Likelihoods[name]
		if ok {
			m.SigMask |= 1 << uint(s)
		}
		for d, dn := range gpu.Domains {
			if p, has := ll[dn]; ok && has {
				m.Lik[s][d] = p
			} else {
				m.Lik[s][d] = math.NaN() // likelihoodFor: 0.5 (bayesian.go:293-296)
			}
		}
	}
	// rows of Likelihoods outside the twelve known signals have no threshold (bayesian.go:222), are never elevated,
	// and add the constant log clamp(1 - p) per domain
	for name, ll := range b.Likelihoods {
		if gpu.SignalIndex(name) != gpu.SigUnsupported {
			continue
		}
		for d, dn := range gpu.Domains {
			p := 0.5
			if v, has := ll[dn]; has {
				p = clampLikelihood(1 - v)
			}
			m.LogExtra[d] += math.Log(p)
		}
	}
	return m
}

// lastModel is the model the library's default context holds (guarded by the context's own mutex: run is only
// entered through it).
var (
	lastModel    gpu.Model
	lastModelSet bool
)

func (b *BayesianAttributor) run(signals []map[string]float64, ruleLabel []uint8) ([]float64, []uint8, []uint16) {
	ctx := gpu.MustDefault()
	// the flattened model goes to the library only when Priors / Likelihoods changed since the last call
	if m := b.model(); !lastModelSet || m != lastModel {
		if err := ctx.SetModel(&m); err != nil {
			panic(err)
		}
		lastModel, lastModelSet = m, true
	}
	n := len(signals)
	arr := make([]float64, n*gpu.NSig)
	has := make([]uint8, n)
	for i := range arr {
		arr[i] = math.NaN()
	}
	for i, sm := range signals {
		if len(sm) > 0 {
			has[i] = 1
		}
		for k, v := range sm {
			if j := gpu.SignalIndex(k); j != gpu.SigUnsupported {
				arr[i*gpu.NSig+int(j)] = v
			}
		}
	}
	post, top1, elev, err := ctx.Attribute(arr, has, ruleLabel)
	if err != nil {
		panic(err)
	}
	return post, top1, elev
}

// posteriors rebuilds []Posterior from one row: evidence = elevated signals whose likelihood for the domain is
// >= 0.5, sorted (bayesian.go:262-276); stable sort by posterior descending (bayesian.go:279-281).
func (b *BayesianAttributor) posteriors(post []float64, elev uint16) []Posterior {
	out := make([]Posterior, 0, gpu.NDom)
	for d, dn := range gpu.Domains {
		var ev []string
		for s, name := range gpu.Signals {
			if elev>>uint(s)&1 == 1 {
				if ll, ok := b.Likelihoods[name]; ok && ll[dn] >= 0.5 {
					ev = append(ev, name)
				}
			}
		}
		sort.Strings(ev)
		out = append(out, Posterior{Domain: dn, Posterior: post[d], Evidence: ev})
	}
	sort.SliceStable(out, func(i, j int) bool { return out[i].Posterior > out[j].Posterior })
	return out
}

// Attribute computes Bayesian posteriors over fault domains given observed signals (GPU build of bayesian.go:218-284).
func (b *BayesianAttributor) Attribute(signals map[string]float64) []Posterior {
	// Attribute itself has no empty-map shortcut: force has_signals by passing a non-empty marker row
	post, _, elev := b.run([]map[string]float64{withMarker(signals)}, nil)
	return b.posteriors(post[:gpu.NDom], elev[0])
}

// withMarker returns signals unchanged when non-empty; for an empty map a map with one unknown key, which the
// packer ignores but which makes len() > 0 (no signal elevated, every domain gets its "not elevated" product).
func withMarker(signals map[string]float64) map[string]float64 {
	if len(signals) > 0 {
		return signals
	}
	return map[string]float64{"": 0}
}

// BuildAttributionsBayes converts samples into attribution envelopes using Bayesian multi-fault inference (GPU build
// of pipeline.go:31-40 + AttributeSample, bayesian.go:315-343): one kernel launch for all samples.
func BuildAttributionsBayes(samples []FaultSample, attributor *BayesianAttributor) []schema.IncidentAttribution {
	if attributor == nil {
		attributor = NewBayesianAttributor()
	}
	out := make([]schema.IncidentAttribution, 0, len(samples))
	if len(samples) == 0 {
		return out
	}
	sig := make([]map[string]float64, len(samples))
	rule := make([]uint8, len(samples))
	for i, s := range samples {
		sig[i] = s.Signals
		rule[i] = uint8(domainIndex(MapFaultLabel(s.FaultLabel)))
	}
	post, top1, elev := attributor.run(sig, rule)
	for i, s := range samples {
		base := BuildAttribution(s)
		if len(s.Signals) == 0 {
			out = append(out, base)
			continue
		}
		row := post[i*gpu.NDom : (i+1)*gpu.NDom]
		ps := attributor.posteriors(row, elev[i])
		hyp := make([]schema.FaultHypothesis, 0, len(ps))
		for _, p := range ps {
			if p.Posterior < 0.01 {
				continue
			}
			hyp = append(hyp, schema.FaultHypothesis{Domain: p.Domain, Posterior: p.Posterior, Evidence: p.Evidence})
		}
		base.FaultHypotheses = hyp
		base.PredictedFaultDomain = gpu.Domains[top1[i]]
		base.Confidence = row[top1[i]]
		out = append(out, base)
	}
	return out
}

// AttributeSample runs Bayesian attribution on one FaultSample (bayesian.go:315-343).
func (b *BayesianAttributor) AttributeSample(sample FaultSample) schema.IncidentAttribution {
	return BuildAttributionsBayes([]FaultSample{sample}, b)[0]
}

type encoded struct {
	n                  int
	actual, pred, expN []uint8
	exp                []uint32
	post               []float64
	names              [16]string
}

// encode turns samples + predictions into the integer columns lsg_confusion takes.
func encode(samples []FaultSample, predictions []schema.IncidentAttribution) encoded {
	n := len(samples)
	if len(predictions) < n {
		n = len(predictions)
	}
	e := encoded{n: n, actual: make([]uint8, n), pred: make([]uint8, n), expN: make([]uint8, n), exp: make([]uint32, n),
		post: make([]float64, n*gpu.NDom)}
	copy(e.names[:], gpu.Domains[:])
	for i := range e.post {
		e.post[i] = math.NaN()
	}
	other := map[string]uint8{}
	for i := 0; i < n; i++ {
		s, p := samples[i], predictions[i]
		a := s.ExpectedDomain
		if a == "" {
			a = MapFaultLabel(s.FaultLabel)
		}
		if d := domainIndex(a); d >= 0 {
			e.actual[i] = uint8(d)
		} else {
			c, ok := other[a]
			if !ok {
				if len(other) >= gpu.MaxActual-gpu.NDom {
					panic("more than 8 actual classes outside AllDomains") // split the batch
				}
				c = uint8(gpu.NDom + len(other))
				other[a] = c
				e.names[c] = a
			}
			e.actual[i] = c
		}
		e.pred[i] = uint8(domainIndex(p.PredictedFaultDomain))
		expected := s.ExpectedDomains
		if len(expected) == 0 {
			expected = []string{a}
		}
		if len(expected) > 8 {
			panic("more than 8 expected domains")
		}
		for k, name := range expected {
			code := uint32(0xF)
			if d := domainIndex(name); d >= 0 {
				code = uint32(d)
			}
			e.exp[i] |= code << (4 * uint(k))
		}
		e.expN[i] = uint8(len(expected))
		for _, h := range p.FaultHypotheses {
			if d := domainIndex(h.Domain); d >= 0 {
				e.post[i*gpu.NDom+d] = h.Posterior
			}
		}
	}
	return e
}

// One round trip serves BuildConfusionMatrix, Accuracy, PartialAccuracy and CoverageAccuracy when they are called on
// the same slices one after the other, as cmd/attributor (main.go:258-308) and pkg/benchmark (harness.go:85-113) do:
// the last result is kept, keyed by the identity of the two slices and the coverage threshold (the matrix, the
// accuracy and the partial numerators do not depend on the threshold).
type confusionKey struct {
	s0        *FaultSample
	p0        *schema.IncidentAttribution
	ns, np    int
	threshold float64
}

var (
	lastConfKey confusionKey
	lastConfOut gpu.Confusion
	lastConfEnc encoded
	lastConfOK  bool
)

func confusion(samples []FaultSample, predictions []schema.IncidentAttribution, threshold float64, needCoverage bool) (gpu.Confusion, encoded) {
	key := confusionKey{&samples[0], &predictions[0], len(samples), len(predictions), threshold}
	if lastConfOK && key.s0 == lastConfKey.s0 && key.p0 == lastConfKey.p0 && key.ns == lastConfKey.ns && key.np == lastConfKey.np &&
		(!needCoverage || key.threshold == lastConfKey.threshold) {
		return lastConfOut, lastConfEnc
	}
	e := encode(samples, predictions)
	out, err := gpu.MustDefault().ConfusionCounts(e.actual, e.pred, e.exp, e.expN, e.post, threshold, 0)
	if err != nil {
		panic(err)
	}
	lastConfKey, lastConfOut, lastConfEnc, lastConfOK = key, out, e, true
	return out, e
}

// InvalidateConfusionCache drops the kept result; call it after mutating samples or predictions in place.
func InvalidateConfusionCache() { lastConfOK = false }

// BuildConfusionMatrix returns matrix counts keyed by actual/predicted fault domain (GPU build of pipeline.go:65-82).
func BuildConfusionMatrix(samples []FaultSample, predictions []schema.IncidentAttribution) map[MatrixKey]int {
	matrix := map[MatrixKey]int{}
	if len(samples) == 0 || len(predictions) == 0 {
		return matrix
	}
	out, e := confusion(samples, predictions, 0, false)
	for a := 0; a < gpu.MaxActual; a++ {
		for p := 0; p < gpu.NDom; p++ {
			if c := out.Matrix[a][p]; c != 0 {
				matrix[MatrixKey{Actual: e.names[a], Predicted: gpu.Domains[p]}] = int(c)
			}
		}
	}
	return matrix
}

// Accuracy returns exact-match ratio for predicted vs expected fault domain (pipeline.go:85-104).
func Accuracy(samples []FaultSample, predictions []schema.IncidentAttribution) float64 {
	if len(predictions) == 0 || len(samples) == 0 {
		return 0
	}
	out, _ := confusion(samples, predictions, 0, false)
	return float64(out.Correct) / float64(len(predictions))
}

// PartialAccuracy: top-1 prediction is one of the expected domains (pipeline.go:110-136).
func PartialAccuracy(samples []FaultSample, predictions []schema.IncidentAttribution) float64 {
	if len(predictions) == 0 || len(samples) == 0 {
		return 0
	}
	out, _ := confusion(samples, predictions, 0, false)
	return float64(out.Partial) / float64(len(predictions))
}

// CoverageAccuracy: mean over incidents of |expected ∩ (hypotheses >= threshold ∪ top-1)| / |expected|
// (pipeline.go:140-185), formed from the library's (found, len) histogram.
func CoverageAccuracy(samples []FaultSample, predictions []schema.IncidentAttribution, threshold float64) float64 {
	if len(predictions) == 0 || len(samples) == 0 {
		return 0
	}
	out, e := confusion(samples, predictions, threshold, true)
	total := 0.0
	for f := 0; f <= gpu.NDom; f++ {
		for ln := 1; ln <= gpu.NDom; ln++ {
			if c := out.Coverage[f][ln]; c != 0 {
				total += float64(c) * (float64(f) / float64(ln))
			}
		}
	}
	if e.n == 0 {
		return 0
	}
	return total / float64(e.n)
}
