//go:build gpu

package gpu

import (
	"fmt"
	"time"
)

// Interner turns the reference's string-keyed records into packed records: five dictionaries (0 = empty string),
// one identity row per distinct (pod, node, service), and the join group the C ABI's contract asks for — union-find
// over rows that share a pod (tiers 2-3, dns.go:61-68), a (service, node) pair (tier 4, dns.go:69-73) or carry the
// same trace id (tier 1, dns.go:58-60).  One interner per batch: spans and signals must share ids.
type Interner struct {
	trace, pod, node, service, conn map[string]uint32
	rows                            map[[3]uint32]uint32
	rowList                         [][3]uint32
	parent                          []uint32
	byPod, byTrace                  map[uint32]uint32
	bySvcNode                       map[[2]uint32]uint32
}

func NewInterner() *Interner {
	d := func() map[string]uint32 { return map[string]uint32{"": 0} }
	return &Interner{trace: d(), pod: d(), node: d(), service: d(), conn: d(), rows: map[[3]uint32]uint32{},
		byPod: map[uint32]uint32{}, byTrace: map[uint32]uint32{}, bySvcNode: map[[2]uint32]uint32{}}
}

func id(d map[string]uint32, s string) uint32 {
	if v, ok := d[s]; ok {
		return v
	}
	v := uint32(len(d))
	d[s] = v
	return v
}

func (in *Interner) find(x uint32) uint32 {
	for in.parent[x] != x {
		in.parent[x] = in.parent[in.parent[x]]
		x = in.parent[x]
	}
	return x
}

func (in *Interner) union(a, b uint32) {
	ra, rb := in.find(a), in.find(b)
	if ra == rb {
		return
	}
	if ra < rb {
		in.parent[rb] = ra
	} else {
		in.parent[ra] = rb
	}
}

func (in *Interner) row(pod, node, service string) uint32 {
	key := [3]uint32{id(in.pod, pod), id(in.node, node), id(in.service, service)}
	if r, ok := in.rows[key]; ok {
		return r
	}
	r := uint32(len(in.rowList))
	in.rows[key] = r
	in.rowList = append(in.rowList, key)
	in.parent = append(in.parent, r)
	if key[0] != 0 {
		if other, ok := in.byPod[key[0]]; ok {
			in.union(other, r)
		} else {
			in.byPod[key[0]] = r
		}
	}
	if key[1] != 0 && key[2] != 0 {
		sn := [2]uint32{key[2], key[1]}
		if other, ok := in.bySvcNode[sn]; ok {
			in.union(other, r)
		} else {
			in.bySvcNode[sn] = r
		}
	}
	return r
}

func (in *Interner) traceID(trace string, row uint32) (uint32, error) {
	t := id(in.trace, trace)
	if t >= 1<<TraceBits {
		return 0, fmt.Errorf("more than 2^28 distinct trace ids in one batch")
	}
	if t != 0 {
		if other, ok := in.byTrace[t]; ok {
			in.union(other, row)
		} else {
			in.byTrace[t] = row
		}
	}
	return t, nil
}

func unixNS(ts time.Time) int64 {
	if ts.IsZero() {
		return TsZero
	}
	return ts.UnixNano()
}

func pid(p int) uint32 {
	if p <= 0 { // PID > 0 guard, dns.go:62
		return 0
	}
	return uint32(p)
}

// Span packs one SpanRef / SpanRecord.  incident is the incident index of the span or None.
func (in *Interner) Span(traceID, service, node, pod string, p int, conn string, ts time.Time, incident uint32) (Span, error) {
	r := in.row(pod, node, service)
	t, err := in.traceID(traceID, r)
	return Span{TsNs: unixNS(ts), Ident: r, PID: pid(p), Conn: id(in.conn, conn), Trace: t, Incident: incident}, err
}

// SignalIndex is the slot of a signal name, or SigUnsupported (signalAttrKey's default branch, correlator.go:169-171).
func SignalIndex(name string) uint32 {
	for i, s := range Signals {
		if s == name {
			return uint32(i)
		}
	}
	return SigUnsupported
}

// Event packs one SignalRef.
func (in *Interner) Event(signal, traceID, service, node, pod string, p int, conn string, ts time.Time, value float64) (Event, error) {
	r := in.row(pod, node, service)
	t, err := in.traceID(traceID, r)
	return Event{TsNs: unixNS(ts), Value: value, Ident: r, PID: pid(p), Conn: id(in.conn, conn),
		TraceSig: t<<4 | SignalIndex(signal)}, err
}

// Idents is the identity table with join groups: 1 + index of the row's union-find component.  Call it after
// every span and event of the batch has been packed.
func (in *Interner) Idents() []Ident {
	out := make([]Ident, len(in.rowList))
	comp := map[uint32]uint32{}
	for r, key := range in.rowList {
		root := in.find(uint32(r))
		g, ok := comp[root]
		if !ok {
			g = uint32(len(comp)) + 1
			comp[root] = g
		}
		out[r] = Ident{Pod: key[0], Node: key[1], Service: key[2], Grp: g}
	}
	if len(out) == 0 {
		out = append(out, Ident{})
	}
	return out
}
