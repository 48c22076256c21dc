package foremastoracle

// Checks the Go transcription against the fixtures the C oracle and the CUDA path are held to:
// tests/golden/rank_cases.json (41 seeded cases, scipy cross-checked) and
// tests/golden/demo_fixture.json (the reference's examples/spring-boot-demo data1/data2).
// Run from oracle/go:  go test ./...

import (
	"encoding/json"
	"math"
	"os"
	"path/filepath"
	"testing"
)

const relTol = 1e-5 // the bar of BASELINE.json for p-values and smoothed values

type rankCase struct {
	X, Y       []float64
	Degenerate bool    `json:"degenerate"`
	MW2R1      int64   `json:"mw_2r1"`
	MW2U1      int64   `json:"mw_2u1"`
	MWTies     int64   `json:"mw_ties"`
	PMW        float64 `json:"p_mw"`
	KWH        float64 `json:"kw_h"`
	PK         float64 `json:"p_k"`
	WN         uint32  `json:"w_n"`
	W2WPlus    int64   `json:"w_2wplus"`
	W2WMinus   int64   `json:"w_2wminus"`
	WTies      int64   `json:"w_ties"`
	PW         float64 `json:"p_w"`
}

func golden(t *testing.T, name string, v interface{}) {
	t.Helper()
	raw, err := os.ReadFile(filepath.Join("..", "..", "..", "tests", "golden", name))
	if err != nil {
		t.Fatal(err)
	}
	if err := json.Unmarshal(raw, v); err != nil {
		t.Fatal(err)
	}
}

func f32(v []float64) []float32 {
	out := make([]float32, len(v))
	for i, x := range v {
		out[i] = float32(x)
	}
	return out
}

func near(a, b float64) bool { return math.Abs(a-b) <= relTol*math.Abs(b) }

func rankParams() Params {
	p := DefaultParams()
	p.MinMannWhite, p.MinWilcoxon, p.MinKruskal = 1, 1, 1
	return p
}

func TestRankCases(t *testing.T) {
	var doc struct {
		Cases []rankCase `json:"cases"`
	}
	golden(t, "rank_cases.json", &doc)
	if len(doc.Cases) == 0 {
		t.Fatal("no cases")
	}
	p := rankParams()
	for i, c := range doc.Cases {
		s := ScoreUnit(&p, nil, f32(c.X), f32(c.Y), 0, nil, nil)
		if c.Degenerate {
			if s.PMW != 1.0 || s.PK != 1.0 || s.PW != 1.0 {
				t.Errorf("case %d: degenerate case must give p = 1", i)
			}
			continue
		}
		if s.MW2R1 != c.MW2R1 || s.MW2U1 != c.MW2U1 || s.MWTies != c.MWTies || s.WN != c.WN {
			t.Errorf("case %d: rank sums %d %d %d %d, want %d %d %d %d", i, s.MW2R1, s.MW2U1, s.MWTies, s.WN,
				c.MW2R1, c.MW2U1, c.MWTies, c.WN)
		}
		if c.WN > 0 && (s.W2WPlus != c.W2WPlus || s.W2WMinus != c.W2WMinus || s.WTies != c.WTies) {
			t.Errorf("case %d: signed-rank sums %d %d %d, want %d %d %d", i, s.W2WPlus, s.W2WMinus, s.WTies,
				c.W2WPlus, c.W2WMinus, c.WTies)
		}
		if !near(s.PMW, c.PMW) || !near(s.PK, c.PK) {
			t.Errorf("case %d: p_mw %g p_k %g, want %g %g", i, s.PMW, s.PK, c.PMW, c.PK)
		}
	}
}

func TestDemoFixture(t *testing.T) {
	var doc struct {
		Data1  []float64 `json:"data1"`
		Data2  []float64 `json:"data2"`
		Pooled rankCase  `json:"pooled"`
		MA     struct {
			Mean      float64  `json:"mean"`
			Std       float64  `json:"std"`
			UpperK2   float64  `json:"upper_k2"`
			Anomalies []uint32 `json:"anomalies_idx"`
		} `json:"ma"`
	}
	golden(t, "demo_fixture.json", &doc)
	p := rankParams()
	up, lo := make([]float32, len(doc.Data2)), make([]float32, len(doc.Data2))
	p.BandAlgorithm = AlgMAAll
	// data1 is both history and baseline, data2 the current window (DECISIONS D15)
	s := ScoreUnit(&p, f32(doc.Data1), f32(doc.Data1), f32(doc.Data2), MtOther, up, lo)
	c := doc.Pooled
	if s.MW2R1 != c.MW2R1 || s.MW2U1 != c.MW2U1 || s.MWTies != c.MWTies {
		t.Errorf("pooled rank sums %d %d %d, want %d %d %d", s.MW2R1, s.MW2U1, s.MWTies, c.MW2R1, c.MW2U1, c.MWTies)
	}
	if s.WN != c.WN || s.W2WPlus != c.W2WPlus || s.W2WMinus != c.W2WMinus || s.WTies != c.WTies {
		t.Errorf("signed-rank sums %d %d %d %d", s.WN, s.W2WPlus, s.W2WMinus, s.WTies)
	}
	if !near(s.PMW, c.PMW) || !near(s.PK, c.PK) || !near(s.PW, c.PW) || !near(s.KWH, c.KWH) {
		t.Errorf("p-values %g %g %g H %g", s.PMW, s.PK, s.PW, s.KWH)
	}
	if s.Status&SDiffers != 0 {
		t.Error("the three tests must say: same distribution")
	}
	if !near(s.MAMean, doc.MA.Mean) || !near(s.MAStd, doc.MA.Std) || !near(float64(up[0]), doc.MA.UpperK2) {
		t.Errorf("band %g %g %g", s.MAMean, s.MAStd, up[0])
	}
	if s.NAnom[0] != uint32(len(doc.MA.Anomalies)) || s.FirstIdx[0] != doc.MA.Anomalies[0] {
		t.Errorf("anomalies %d first %d, want %v", s.NAnom[0], s.FirstIdx[0], doc.MA.Anomalies)
	}
	// ... and the service is unhealthy although the distributions agree (design.md:35, :43)
	units := []Unit{{Hist: f32(doc.Data1), Base: f32(doc.Data1), Cur: f32(doc.Data2)}}
	v := Verdicts(units, []SeriesScore{s}, 1, 0)
	if v[0].Verdict != VUnhealth || v[0].WorstUnit != 0 {
		t.Errorf("verdict %+v", v[0])
	}
}
