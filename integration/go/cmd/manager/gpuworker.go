// +build gpu

// gpuworker.go -- the in-process scoring worker that takes the brain's place at the Elasticsearch seam of
// foremast-service.  NOT COMPILED in the repository that ships it (no Go toolchain in its build image);
// foremast_b200/brain.py::BrainWorker is the same loop in Python and is what that repository's tests drive.
//
// Wiring (cmd/manager/main.go): start `go gpuWorkerLoop(elasticClient)` next to the router; RegisterEntry
// keeps writing documents with Status "initial" (pkg/search/elasticsearchstore.go:58-87) and SearchByID keeps
// serving them (main.go:227-255) -- both untouched.
package main

import (
	"encoding/json"
	"fmt"
	"html"
	"io/ioutil"
	"net/http"
	"strings"
	"time"

	"foremast.ai/foremast/foremast-service/pkg/gpu"
	"foremast.ai/foremast/foremast-service/pkg/models"
)

const (
	maxStuckInSeconds = 90 // deploy/foremast/3_brain/foremast-brain.yaml:80-81
	pollEvery         = 10 * time.Second
)

// JobStore is what the worker needs from pkg/search: documents by status (ByStatus,
// elasticsearchstore.go:154-180, "used by backend python model") and a partial update.
type JobStore interface {
	ByStatus(statuses []string, limit int) ([]models.DocumentResponse, error)
	Update(id string, fields map[string]interface{}) error
}

// splitConfig: "alias== url || alias== url" (convertMetricQuerys, main.go:50-81; separators main.go:29,32).
func splitConfig(cfg string) map[string]string {
	out := map[string]string{}
	for _, part := range strings.Split(cfg, " ||") {
		kv := strings.SplitN(strings.TrimSpace(part), "== ", 2)
		if len(kv) == 2 {
			out[strings.TrimSpace(kv[0])] = strings.TrimSpace(kv[1])
		}
	}
	return out
}

func fetch(url string, start, end int64) ([]byte, error) {
	// continuous / hpa jobs carry placeholders (main.go:59-63)
	url = strings.Replace(strings.Replace(url, "START_TIME", fmt.Sprint(start), -1), "END_TIME", fmt.Sprint(end), -1)
	resp, err := http.Get(url)
	if err != nil {
		return nil, err
	}
	defer resp.Body.Close()
	return ioutil.ReadAll(resp.Body)
}

func metricType(alias string) uint8 {
	a := strings.ToLower(alias)
	switch {
	case strings.Contains(a, "5xx"):
		return 1
	case strings.Contains(a, "4xx"):
		return 2
	case strings.Contains(a, "latency"):
		return 3
	case strings.Contains(a, "cpu"):
		return 4
	case strings.Contains(a, "mem"):
		return 5
	}
	return 0
}

// one cycle: claim -> fetch -> ONE scoring pass for all claimed documents -> write back
func gpuWorkerCycle(ctx *gpu.Ctx, params *gpu.Params, store JobStore) error {
	now := time.Now().Unix()
	docs, err := store.ByStatus([]string{"initial", "postprocess_inprogress", "preprocess_inprogress", "preprocess_completed"}, 0)
	if err != nil {
		return err
	}
	var claimed []models.DocumentResponse
	for _, d := range docs {
		stuck := d.Status == "preprocess_inprogress" || d.Status == "preprocess_completed"
		if stuck && d.ModifiedAt != nil && now-d.ModifiedAt.Unix() <= maxStuckInSeconds {
			continue // another worker is on it (docs/guides/design.md:39)
		}
		store.Update(d.ID, map[string]interface{}{"status": "preprocess_inprogress", "modified_at": time.Now()})
		claimed = append(claimed, d)
	}
	var jobs []gpu.Job
	failed := map[int]string{}
	for i, d := range claimed {
		start, end := d.StartTimeUnix(), d.EndTimeUnix()
		cur, base, hist := splitConfig(d.CurrentConfig), splitConfig(d.BaselineConfig), splitConfig(d.HistoricalConfig)
		for alias, url := range cur {
			j := gpu.Job{ServiceID: uint32(i), MetricType: metricType(alias), Step: 60}
			if j.Cur, err = fetch(url, start, end); err != nil {
				failed[i] = err.Error()
				break
			}
			if u, ok := base[alias]; ok {
				j.Base, _ = fetch(u, start, end)
			}
			if u, ok := hist[alias]; ok {
				j.Hist, _ = fetch(u, now-7*86400, now)
			}
			jobs = append(jobs, j)
		}
		store.Update(d.ID, map[string]interface{}{"status": "preprocess_completed", "modified_at": time.Now()})
	}
	verdicts, _, err := ctx.ScoreJobs(jobs, params, len(claimed))
	if err != nil {
		return err
	}
	for i, d := range claimed {
		fields := map[string]interface{}{"statusCode": "200", "modified_at": time.Now()}
		switch {
		case failed[i] != "":
			fields["status"], fields["statusCode"], fields["reason"] = "preprocess_failed", "500", failed[i]
		case verdicts[i].Verdict == gpu.Unhealth:
			fields["status"] = "completed_unhealth"
			// the shape foremast-trigger's regexes read (trigger.go:295,309): HTML-escaped JSON
			body, _ := json.Marshal(map[string]interface{}{"anomaly": []map[string]interface{}{{"name": d.AppName, "ts": []float64{float64(now)}}}})
			fields["reason"] = html.EscapeString(string(body))
		case verdicts[i].Verdict == gpu.Unknown:
			fields["status"], fields["reason"] = "completed_unknown", "current metric is not available"
		case now < d.EndTimeUnix():
			fields["status"] = "postprocess_inprogress" // keep watching until endTime (docs/guides/design.md:43)
		default:
			fields["status"] = "completed_health"
		}
		store.Update(d.ID, fields)
	}
	return nil
}

func gpuWorkerLoop(store JobStore) {
	ctx, err := gpu.Init(-1)
	if err != nil {
		panic(err) // there is no CPU scoring path
	}
	defer ctx.Close()
	params := gpu.DefaultParams()
	for range time.Tick(pollEvery) {
		if err := gpuWorkerCycle(ctx, &params, store); err != nil {
			fmt.Println("gpu worker:", err)
		}
	}
}
