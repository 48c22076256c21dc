// Package foremastoracle is the Go transcription of the CPU oracle (oracle/fm_oracle.c), for
// maintainers of foremast-service who want to check libforemast_gpu from their own test suite
// (BASELINE.json north_star: "a committed CPU Go transcription of the documented algorithms and
// thresholds").  TEST INFRASTRUCTURE ONLY, PARITY UNPINNED -- see oracle/fm_oracle.h and
// oracle/DECISIONS.md, which is the specification both transcriptions follow line by line.
//
// NOT COMPILED IN THE BUILD CONTAINER (it has no Go toolchain); the C oracle is the one the
// test-suite and the benchmark run.  `go test ./...` in this directory checks it against the
// same golden fixtures (tests/golden/).
//
// Numerics contract (DECISIONS "Exact arithmetic"): sequential float64 sums in sample order,
// fused multiply-adds only where the C source says fma() -- written math.FMA here; every other
// product that feeds an addition is wrapped in an explicit float64(...) conversion, which the Go
// specification defines as a rounding point, so that no architecture fuses it.
package foremastoracle

import (
	"math"
	"sort"
)

// ML_ALGORITHM bits (deploy/foremast/3_brain/foremast-brain.yaml:24-25, docs/guides/design.md:68-71).
const (
	AlgMAAll     uint32 = 1 << 0
	AlgMARolling uint32 = 1 << 1
	AlgEWMA      uint32 = 1 << 2
	AlgDES       uint32 = 1 << 3
	AlgHW        uint32 = 1 << 4
	NAlg                = 5
)

// ML_PAIRWISE_ALGORITHM (foremast-brain/README.md:34).
const (
	PwNone uint32 = iota
	PwAll
	PwAny
	PwMannWhite
	PwWilcoxon
	PwKruskal
)

// ML_BOUND as a bit mask (DECISIONS D1).
const (
	BoundUpper uint32 = 1
	BoundLower uint32 = 2
	BoundBoth  uint32 = 3
)

// metric_type rows of foremast-brain.yaml:34-73.
const (
	MtOther = iota
	MtError5xx
	MtError4xx
	MtLatency
	MtCPU
	MtMemory
	MtTPS
	MtGC
	MtThreads
	MaxMetricTypes = 16
)

// Terminal statuses of foremast-service/pkg/converter/converter.go:16-21, "worse" is larger.
const (
	VHealth   uint32 = 0 // completed_health
	VUnknown  uint32 = 1 // completed_unknown
	VUnhealth uint32 = 2 // completed_unhealth
)

// Per-unit status bits (fm_series_score.status).
const (
	SUnknown     uint32 = 1 << 0
	SAnomaly     uint32 = 1 << 1
	SDiffers     uint32 = 1 << 2
	SMWRun       uint32 = 1 << 3
	SWilcoxonRun uint32 = 1 << 4
	SKruskalRun  uint32 = 1 << 5
	SModelBase   uint32 = 1 << 6
	SAlgShift           = 8
)

// NoneIdx marks "no anomalous point" in FirstIdx.
const NoneIdx uint32 = 0xFFFFFFFF

// MetricRule is one row of the threshold table, foremast-brain.yaml:26-73.
// RankMaxSegment is FM_RANK_MAX_SEGMENT of include/foremast_gpu.h (DECISIONS D17).
const RankMaxSegment = 2048

type MetricRule struct {
	Threshold     float64
	MinLowerBound float64
	Bound         uint32
}

// Params mirrors fm_params (include/foremast_gpu.h).
type Params struct {
	Algorithms        uint32
	Pairwise          uint32
	PairwiseThreshold float64
	DifferScale       float64
	MinMannWhite      uint32
	MinWilcoxon       uint32
	MinKruskal        uint32
	MinHistorical     uint32
	MinAnomalyPoints  uint32
	MAWindow          uint32
	Period            uint32
	BandAlgorithm     uint32
	Alpha, Beta, Gamma float64
	Rules             [MaxMetricTypes]MetricRule
}

// SeriesScore mirrors fm_series_score.
type SeriesScore struct {
	Status   uint32
	NAnom    [NAlg]uint32
	FirstIdx [NAlg]uint32
	WN       uint32
	MW2U1    int64
	MW2R1    int64
	MWTies   int64
	W2WPlus  int64
	W2WMinus int64
	WTies    int64
	PMW, PW, PK float64
	KWH      float64
	KEff     float64
	MAMean   float64
	MAStd    float64
	Sigma    [NAlg]float64
}

// Verdict mirrors fm_verdict.
type Verdict struct {
	ServiceID       uint32
	Verdict         uint32
	NUnits          uint32
	NUnknown        uint32
	NAnomUnits      uint32
	NAnomPoints     uint32
	WorstUnit       uint32
	WorstUnitPoints uint32
}

// Unit is one series-window: historical | baseline | current samples of one job alias
// (historicalConfig / baselineConfig / currentConfig, foremast-service/pkg/models/models.go:108-110).
type Unit struct {
	Hist, Base, Cur []float32
	ServiceID       uint32
	MetricType      uint8
}

// DefaultParams: documented defaults of foremast-brain.yaml:24-79 plus DECISIONS D2, D5, D6, D10.
func DefaultParams() Params {
	var p Params
	p.Algorithms = AlgMAAll // ML_ALGORITHM=moving_average_all, brain.yaml:24-25
	p.Pairwise = PwAll
	p.PairwiseThreshold = 0.05
	p.DifferScale = 0.5
	p.MinMannWhite = 20 // brain.yaml:74-75
	p.MinWilcoxon = 20  // brain.yaml:76-77
	p.MinKruskal = 5    // brain.yaml:78-79
	p.MinHistorical = 10
	p.MinAnomalyPoints = 1
	p.MAWindow = 30
	p.Period = 1440
	p.Alpha, p.Beta, p.Gamma = 0.3, 0.1, 0.1
	// brain.yaml:26-73: threshold / bound / min_lower_bound rows
	p.Rules[MtOther] = MetricRule{2.0, 0.0, BoundUpper}
	p.Rules[MtError5xx] = MetricRule{2.0, 0.0, BoundUpper}
	p.Rules[MtError4xx] = MetricRule{3.0, 0.0, BoundUpper}
	p.Rules[MtLatency] = MetricRule{10.0, 0.0, BoundBoth}
	p.Rules[MtCPU] = MetricRule{5.0, 0.0, BoundUpper}
	p.Rules[MtMemory] = MetricRule{5.0, 0.0, BoundUpper}
	return p
}

func ruleFor(p *Params, mt uint8) *MetricRule {
	i := 0
	if int(mt) < MaxMetricTypes {
		i = int(mt)
	}
	if p.Rules[i].Bound == 0 { // D16: an unset row falls back to row 0
		i = 0
	}
	return &p.Rules[i]
}

// ---------------------------------------------------------------- rank statistics
// Mann-Whitney / Wilcoxon / Kruskal-Wallis, foremast-brain/README.md:34-38; design.md:91.

// fkey is the order-preserving map of a finite float32 onto uint32.
func fkey(f float32) uint32 {
	u := math.Float32bits(f)
	if u&0x80000000 != 0 {
		return ^u
	}
	return u | 0x80000000
}

// rankGroups: keys are (value_key << 1 | tag), sorted; groups are runs of equal value_key.
// Returns the sum over groups of cnt_tag0 * (2s + t + 1) and the tie term sum(t^3 - t).
func rankGroups(k []uint64) (r2, ties int64) {
	n := len(k)
	for s := 0; s < n; {
		v := k[s] >> 1
		e, c0 := s, int64(0)
		for e < n && k[e]>>1 == v {
			c0 += int64(^k[e] & 1)
			e++
		}
		t := int64(e - s)
		r2 += c0 * (2*int64(s) + t + 1)
		ties += t*t*t - t
		s = e
	}
	return
}

func pMannWhitney(u1x2 int64, n1, n2 uint32, ties int64) float64 {
	nn := float64(float64(n1) * float64(n2))
	N := float64(n1) + float64(n2)
	v := float64(nn/12.0) * ((N + 1.0) - float64(float64(ties)/float64(N*(N-1.0))))
	if !(v > 0.0) {
		return 1.0
	}
	u2x2 := 2*int64(n1)*int64(n2) - u1x2
	umax := u1x2
	if u2x2 > umax {
		umax = u2x2
	}
	z := (float64(float64(umax-int64(n1)*int64(n2))*0.5) - 0.5) / math.Sqrt(v)
	p := math.Erfc(float64(z * 0.70710678118654752440))
	if p > 1.0 {
		return 1.0
	}
	return p
}

func pKruskal(r1x2 int64, n1, n2 uint32, ties int64) (p, h float64) {
	N := float64(n1) + float64(n2)
	Ni := int64(n1) + int64(n2)
	r2x2 := Ni*(Ni+1) - r1x2
	r1, r2 := float64(0.5*float64(r1x2)), float64(0.5*float64(r2x2))
	c := 1.0 - float64(float64(ties)/(float64(float64(N*N)*N)-N))
	if !(c > 0.0) {
		return 1.0, 0.0
	}
	a := float64(float64(r1*r1) / float64(n1))
	b := float64(float64(r2*r2) / float64(n2))
	h = float64(float64(12.0/float64(N*(N+1.0)))*(a+b)) - float64(3.0*(N+1.0))
	h = h / c
	if !(h > 0.0) {
		h = 0.0
	}
	return math.Erfc(math.Sqrt(float64(h * 0.5))), h
}

func pWilcoxon(wp2, wm2 int64, n uint32, ties int64) float64 {
	if n == 0 {
		return 1.0
	}
	nd := float64(n)
	se2 := float64(float64(float64(nd*(nd+1.0))*(float64(2.0*nd)+1.0))/24.0) - float64(float64(ties)/48.0)
	if !(se2 > 0.0) {
		return 1.0
	}
	tmin := wp2
	if wm2 < tmin {
		tmin = wm2
	}
	z := (float64(float64(tmin)*0.5) - float64(float64(nd*(nd+1.0))/4.0)) / math.Sqrt(se2)
	p := math.Erfc(float64(math.Abs(z) * 0.70710678118654752440))
	if p > 1.0 {
		return 1.0
	}
	return p
}

func sortKeys(k []uint64) { sort.Slice(k, func(i, j int) bool { return k[i] < k[j] }) }

// ---------------------------------------------------------------- band models
// design.md:68-71 (Moving Average, Exponential / Double Exponential Smoothing, Holt-Winters);
// threshold / bound / min_lower_bound semantics README.md:24-25, brain.yaml:26-33.

type bandRule struct {
	k, mlb float64
	bound  uint32
}

func outside(x, up, lo float64, bound uint32) bool {
	return (bound&1 != 0 && x > up) || (bound&2 != 0 && x < lo)
}

func band(centre, sigma float64, r *bandRule) (up, lo float64) {
	up = math.FMA(r.k, sigma, centre)
	l := math.FMA(-r.k, sigma, centre)
	if l < r.mlb {
		l = r.mlb
	}
	return up, l
}

func note(o *SeriesScore, a int, j uint32) {
	if o.NAnom[a] == 0 {
		o.FirstIdx[a] = j
	}
	o.NAnom[a]++
}

// ScoreUnit scores one series-window.  bandUp / bandLo (optional, len(cur) each) receive the
// per-point band of p.BandAlgorithm.
func ScoreUnit(p *Params, hist, base, cur []float32, mt uint8, bandUp, bandLo []float32) SeriesScore {
	var o SeriesScore
	for a := 0; a < NAlg; a++ {
		o.FirstIdx[a] = NoneIdx
	}
	o.PMW, o.PW, o.PK = 1.0, 1.0, 1.0
	nh, nb, nc := uint32(len(hist)), uint32(len(base)), uint32(len(cur))

	// ---- 1. pairwise (README.md:32-38; gated by MIN_*_DATA_POINTS brain.yaml:74-79)
	wantMW := p.Pairwise == PwAll || p.Pairwise == PwAny || p.Pairwise == PwMannWhite
	wantW := p.Pairwise == PwAll || p.Pairwise == PwAny || p.Pairwise == PwWilcoxon
	wantK := p.Pairwise == PwAll || p.Pairwise == PwAny || p.Pairwise == PwKruskal
	m := nb
	if nc < m {
		m = nc
	}
	// D17: rank tests look at segments of at most RankMaxSegment samples (Wilcoxon: pairs)
	shortEnough := nb <= RankMaxSegment && nc <= RankMaxSegment
	runMW := wantMW && nb >= p.MinMannWhite && nc >= p.MinMannWhite && m > 0 && shortEnough
	runW := wantW && m >= p.MinWilcoxon && m > 0 && m <= RankMaxSegment
	runK := wantK && nb >= p.MinKruskal && nc >= p.MinKruskal && m > 0 && shortEnough
	nRun, nRej := 0, 0
	if runMW || runK {
		keys := make([]uint64, 0, nb+nc)
		for _, x := range base {
			keys = append(keys, uint64(fkey(x+0))<<1)
		}
		for _, x := range cur {
			keys = append(keys, uint64(fkey(x+0))<<1|1)
		}
		sortKeys(keys)
		r1x2, ties := rankGroups(keys)
		o.MW2R1 = r1x2
		o.MW2U1 = r1x2 - int64(nb)*(int64(nb)+1)
		o.MWTies = ties
		if runMW {
			o.Status |= SMWRun
			o.PMW = pMannWhitney(o.MW2U1, nb, nc, ties)
			nRun++
			if o.PMW < p.PairwiseThreshold {
				nRej++
			}
		}
		if runK {
			o.Status |= SKruskalRun
			o.PK, o.KWH = pKruskal(r1x2, nb, nc, ties)
			nRun++
			if o.PK < p.PairwiseThreshold {
				nRej++
			}
		}
	}
	if runW {
		keys := make([]uint64, 0, m)
		for i := uint32(0); i < m; i++ {
			d := cur[i] - base[i]
			if d == 0 {
				continue
			}
			mag := math.Float32bits(float32(math.Abs(float64(d))))
			tag := uint64(0)
			if d < 0 {
				tag = 1
			}
			keys = append(keys, uint64(mag)<<1|tag)
		}
		sortKeys(keys)
		n := uint32(len(keys))
		wp2, ties := rankGroups(keys)
		o.WN = n
		o.W2WPlus = wp2
		o.W2WMinus = int64(n)*(int64(n)+1) - wp2
		o.WTies = ties
		o.Status |= SWilcoxonRun
		o.PW = pWilcoxon(wp2, o.W2WMinus, n, ties)
		nRun++
		if o.PW < p.PairwiseThreshold {
			nRej++
		}
	}
	differs := false
	if p.Pairwise == PwAll {
		differs = nRun > 0 && nRej == nRun
	} else if p.Pairwise != PwNone {
		differs = nRej > 0
	}
	if differs {
		o.Status |= SDiffers
	}

	// ---- 2. threshold (design.md:35 "it will be lower threshold")
	mr := ruleFor(p, mt)
	r := bandRule{mr.Threshold, mr.MinLowerBound, mr.Bound}
	if differs {
		r.k = float64(mr.Threshold * p.DifferScale)
	}
	o.KEff = r.k

	// ---- 3. model window (D15)
	var src []float32
	if nh >= p.MinHistorical && nh > 0 {
		src = hist
	} else if nb >= p.MinHistorical && nb > 0 {
		src = base
		o.Status |= SModelBase
	}
	if src == nil || nc == 0 {
		o.Status |= SUnknown
		return o
	}
	ns := uint32(len(src))
	wantBand := uint32(0)
	if bandUp != nil && bandLo != nil {
		wantBand = p.BandAlgorithm
	}

	// ---- MA_ALL and MA_ROLLING share the shifted sums
	x0 := float64(src[0])
	if p.Algorithms&AlgMAAll != 0 {
		s1, s2 := 0.0, 0.0
		for _, x := range src {
			d := float64(x) - x0
			s1 += d
			s2 = math.FMA(d, d, s2)
		}
		m1 := s1 / float64(ns)
		mean := x0 + m1
		v := math.FMA(-m1, m1, s2/float64(ns))
		if !(v > 0.0) {
			v = 0.0
		}
		sd := math.Sqrt(v)
		up, lo := band(mean, sd, &r)
		o.MAMean, o.MAStd, o.Sigma[0] = mean, sd, sd
		o.Status |= 1 << (SAlgShift + 0)
		for j, x := range cur {
			if outside(float64(x), up, lo, r.bound) {
				note(&o, 0, uint32(j))
			}
			if wantBand == AlgMAAll {
				bandUp[j], bandLo[j] = float32(up), float32(lo)
			}
		}
	}
	if p.Algorithms&AlgMARolling != 0 {
		W := p.MAWindow
		if W == 0 {
			W = 1
		}
		P1 := make([]float64, ns+nc+1)
		P2 := make([]float64, ns+nc+1)
		s1, s2 := 0.0, 0.0
		for t := uint32(0); t < ns+nc; t++ {
			var x float32
			if t < ns {
				x = src[t]
			} else {
				x = cur[t-ns]
			}
			d := float64(x) - x0
			s1 += d
			s2 = math.FMA(d, d, s2)
			P1[t+1], P2[t+1] = s1, s2
		}
		o.Status |= 1 << (SAlgShift + 1)
		for j := uint32(0); j < nc; j++ {
			q := ns + j
			cnt := W
			if q < cnt {
				cnt = q
			}
			a1, a2 := P1[q]-P1[q-cnt], P2[q]-P2[q-cnt]
			m1 := a1 / float64(cnt)
			mean := x0 + m1
			v := math.FMA(-m1, m1, a2/float64(cnt))
			if !(v > 0.0) {
				v = 0.0
			}
			sd := math.Sqrt(v)
			up, lo := band(mean, sd, &r)
			o.Sigma[1] = sd
			if outside(float64(cur[j]), up, lo, r.bound) {
				note(&o, 1, j)
			}
			if wantBand == AlgMARolling {
				bandUp[j], bandLo[j] = float32(up), float32(lo)
			}
		}
	}
	// ---- EWMA (design.md:69)
	if p.Algorithms&AlgEWMA != 0 {
		s, sse := float64(src[0]), 0.0
		for t := uint32(1); t < ns; t++ {
			e := float64(src[t]) - s
			sse = math.FMA(e, e, sse)
			s = math.FMA(p.Alpha, e, s)
		}
		sigma := 0.0
		if ns >= 2 {
			sigma = math.Sqrt(sse / float64(ns-1))
		}
		o.Sigma[2] = sigma
		o.Status |= 1 << (SAlgShift + 2)
		for j, xf := range cur {
			x := float64(xf)
			up, lo := band(s, sigma, &r)
			if outside(x, up, lo, r.bound) {
				note(&o, 2, uint32(j))
			}
			if wantBand == AlgEWMA {
				bandUp[j], bandLo[j] = float32(up), float32(lo)
			}
			e := x - s
			s = math.FMA(p.Alpha, e, s)
		}
	}
	// ---- DES, Holt linear (design.md:70)
	if p.Algorithms&AlgDES != 0 && ns >= 2 {
		l, b, sse := float64(src[0]), float64(src[1])-float64(src[0]), 0.0
		ab := float64(p.Alpha * p.Beta)
		for t := uint32(1); t < ns; t++ {
			f := l + b
			e := float64(src[t]) - f
			sse = math.FMA(e, e, sse)
			l = math.FMA(p.Alpha, e, f)
			b = math.FMA(ab, e, b)
		}
		sigma := math.Sqrt(sse / float64(ns-1))
		o.Sigma[3] = sigma
		o.Status |= 1 << (SAlgShift + 3)
		for j, xf := range cur {
			x, f := float64(xf), l+b
			up, lo := band(f, sigma, &r)
			if outside(x, up, lo, r.bound) {
				note(&o, 3, uint32(j))
			}
			if wantBand == AlgDES {
				bandUp[j], bandLo[j] = float32(up), float32(lo)
			}
			e := x - f
			l = math.FMA(p.Alpha, e, f)
			b = math.FMA(ab, e, b)
		}
	}
	// ---- Holt-Winters additive (design.md:71; 7-day history metricsquery.go:95)
	mper := p.Period
	if p.Algorithms&AlgHW != 0 && mper >= 2 && uint64(ns) >= 2*uint64(mper) {
		S := make([]float64, mper)
		a0, a1 := 0.0, 0.0
		for i := uint32(0); i < mper; i++ {
			a0 += float64(src[i])
		}
		for i := mper; i < 2*mper; i++ {
			a1 += float64(src[i])
		}
		L0, M2 := a0/float64(mper), a1/float64(mper)
		l, b, sse := L0, (M2-L0)/float64(mper), 0.0
		ab, g1 := float64(p.Alpha*p.Beta), float64(p.Gamma*(1.0-p.Alpha))
		for i := uint32(0); i < mper; i++ {
			S[i] = float64(src[i]) - L0
		}
		idx := uint32(0) // t mod m, t starts at m
		for t := mper; t < ns; t++ {
			lb := l + b
			f := lb + S[idx]
			e := float64(src[t]) - f
			sse = math.FMA(e, e, sse)
			l = math.FMA(p.Alpha, e, lb)
			b = math.FMA(ab, e, b)
			S[idx] = math.FMA(g1, e, S[idx])
			if idx++; idx == mper {
				idx = 0
			}
		}
		sigma := math.Sqrt(sse / float64(ns-mper))
		o.Sigma[4] = sigma
		o.Status |= 1 << (SAlgShift + 4)
		for j, xf := range cur {
			x, lb := float64(xf), l+b
			f := lb + S[idx]
			up, lo := band(f, sigma, &r)
			// guarded comparison (DECISIONS "HW"): excursions below 1e-9 relative do not count, so
			// that an implementation that re-associates the recurrence (the CUDA scan) flags the
			// same points
			guard := float64(1e-9 * (math.Abs(f) + math.Abs(x)))
			if (r.bound&1 != 0 && x-up > guard) || (r.bound&2 != 0 && lo-x > guard) {
				note(&o, 4, uint32(j))
			}
			if wantBand == AlgHW {
				bandUp[j], bandLo[j] = float32(up), float32(lo)
			}
			e := x - f
			l = math.FMA(p.Alpha, e, lb)
			b = math.FMA(ab, e, b)
			S[idx] = math.FMA(g1, e, S[idx])
			if idx++; idx == mper {
				idx = 0
			}
		}
	}
	// ---- 5. unit status (D12)
	for a := 0; a < NAlg; a++ {
		if o.Status>>(SAlgShift+uint(a))&1 != 0 && o.NAnom[a] >= p.MinAnomalyPoints && o.NAnom[a] > 0 {
			o.Status |= SAnomaly
		}
	}
	return o
}

// Score scores every unit (ingest canonicalisation included: the caller has dropped non-finite
// samples; -0.0 is read as +0.0, DECISIONS D11).
func Score(units []Unit, p *Params) []SeriesScore {
	out := make([]SeriesScore, len(units))
	for i := range units {
		u := &units[i]
		out[i] = ScoreUnit(p, u.Hist, u.Base, u.Cur, u.MetricType, nil, nil)
	}
	return out
}

// Verdicts folds unit scores into per-service verdicts (D13): terminal statuses of
// converter.go:16-21; fail fast design.md:43.  firstUnit is the global index of units[0]
// (multi-GPU shards).
func Verdicts(units []Unit, scores []SeriesScore, nServices uint32, firstUnit uint64) []Verdict {
	out := make([]Verdict, nServices)
	worst := make([]uint64, nServices)
	for s := range out {
		out[s].ServiceID = uint32(s)
	}
	for u := range units {
		s := units[u].ServiceID
		if s >= nServices {
			panic("service id out of range")
		}
		o := &scores[u]
		out[s].NUnits++
		if o.Status&SUnknown != 0 {
			out[s].NUnknown++
		}
		if o.Status&SAnomaly != 0 {
			pts := uint32(0)
			for a := 0; a < NAlg; a++ {
				if o.NAnom[a] > pts {
					pts = o.NAnom[a]
				}
			}
			out[s].NAnomUnits++
			out[s].NAnomPoints += pts
			key := uint64(pts)<<32 | uint64(0xFFFFFFFF-uint32(firstUnit+uint64(u)))
			if key > worst[s] {
				worst[s] = key
			}
		}
	}
	for s := range out {
		if out[s].NAnomUnits != 0 {
			out[s].Verdict = VUnhealth
			out[s].WorstUnit = 0xFFFFFFFF - uint32(worst[s]&0xFFFFFFFF)
			out[s].WorstUnitPoints = uint32(worst[s] >> 32)
		} else {
			if out[s].NUnknown == out[s].NUnits {
				out[s].Verdict = VUnknown
			} else {
				out[s].Verdict = VHealth
			}
			out[s].WorstUnit = NoneIdx
		}
	}
	return out
}
