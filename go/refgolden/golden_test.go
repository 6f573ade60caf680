// golden_test.go -- pins this repository's CPU oracle to the reference implementation.
//
// It belongs in the PRISTINE reference checkout (github.com/NJUPT-ISL/Yoda-Scheduler), package yoda,
// next to the reference's own scheduler.go -- not in the overlay tree:
//
//	cp <repo>/go/refgolden/golden_test.go <reference>/pkg/yoda/zz_golden_test.go
//	cd <reference> && go test ./pkg/yoda -run TestGolden -args -golden <repo>/tests/golden            # assert
//	cd <reference> && go test ./pkg/yoda -run TestGolden -args -golden <repo>/tests/golden -update    # + write ref_*.json
//
// For every fixture (kat_cluster.json: hand-derived known answers; splitmix_cluster.json: a seeded
// SplitMix64 cluster whose expected numbers come from oracle/yoda_oracle.c) it rebuilds the cluster as
// scv.Scv objects and framework.NodeInfos, takes every pod through the reference's OWN code in the order the
// plugin intends (SURVEY.md section 3) --
//
//	filter.PodFitsNumber / PodFitsMemory / PodFitsClock   as Yoda.Filter combines them (scheduler.go:85-91)
//	collection.CollectMaxValues                           (scheduler.go:95-103)
//	score.CalculateScore + filter.Uint64ToInt64           (scheduler.go:117-128)
//	(*Yoda).NormalizeScore                                (scheduler.go:132-157)
//	lowest-index node holding the top normalized score    (the harness convention for selectHost)
//
// -- and asserts feasible set, the six maxima, raw and normalized scores and the chosen node against the
// fixture.  The parser / uint64->int64 / normalize known answers of kat_cluster.json are asserted too.
// With -update, what the reference computed is written to ref_<fixture>, which tests/test_golden_ref.py
// compares with the oracle from then on.
//
// NOT COMPILED in this repository's build image (no Go toolchain): written against k8s.io/kubernetes
// v1.20.0 and the field types the reference's call sites force on scv.Scv / scv.Card (SURVEY.md 8(a) T1).
package yoda

import (
	"context"
	"encoding/json"
	"flag"
	"io/ioutil"
	"path/filepath"
	"reflect"
	"testing"

	v1 "k8s.io/api/core/v1"
	metav1 "k8s.io/apimachinery/pkg/apis/meta/v1"
	"k8s.io/kubernetes/pkg/scheduler/framework"

	scv "github.com/NJUPT-ISL/SCV/api/v1"

	"github.com/NJUPT-ISL/Yoda-Scheduler/pkg/yoda/collection"
	"github.com/NJUPT-ISL/Yoda-Scheduler/pkg/yoda/filter"
	"github.com/NJUPT-ISL/Yoda-Scheduler/pkg/yoda/score"
)

var (
	goldenDir = flag.String("golden", "", "directory holding kat_cluster.json and splitmix_cluster.json (<repo>/tests/golden)")
	update    = flag.Bool("update", false, "also write ref_<fixture>.json with the reference's outputs")
)

type goldenNode struct {
	Name       string              `json:"name"`
	HasScv     bool                `json:"has_scv"`
	CardNumber uint64              `json:"card_number"`
	FreeSum    uint64              `json:"free_sum"`
	TotalSum   uint64              `json:"total_sum"`
	Cards      [][]json.RawMessage `json:"cards"` // [health, free, total, clock, core, bandwidth, power]
	Pods       []map[string]string `json:"pods"`
}

type goldenCase struct {
	Labels   map[string]string `json:"labels"`
	Feasible []string          `json:"feasible"`
	Max6     []uint64          `json:"max6"` // bandwidth, clock, core, free, power, total (collection.MaxValue order)
	Raw      map[string]int64  `json:"raw"`
	Norm     map[string]int64  `json:"norm"`
	Chosen   *string           `json:"chosen"`
}

type goldenDoc struct {
	Provenance  string            `json:"_provenance,omitempty"`
	Nodes       []goldenNode      `json:"nodes"`
	Cases       []goldenCase      `json:"cases"`
	Parser      []json.RawMessage `json:"parser,omitempty"`          // [string, uint64]
	U64ToI64    []json.RawMessage `json:"uint64_to_int64,omitempty"` // [uint64, int64]
	Normalize   []json.RawMessage `json:"normalize,omitempty"`       // [[in...], [out...]]
}

func u64(t *testing.T, m json.RawMessage) uint64 {
	var v uint64
	if err := json.Unmarshal(m, &v); err != nil {
		t.Fatalf("bad number %s: %v", m, err)
	}
	return v
}

func buildCluster(t *testing.T, doc *goldenDoc) ([]*scv.Scv, scv.ScvList, []*framework.NodeInfo) {
	scvs := make([]*scv.Scv, len(doc.Nodes))
	var list scv.ScvList
	infos := make([]*framework.NodeInfo, len(doc.Nodes))
	for i, nd := range doc.Nodes {
		var running []*v1.Pod
		for k, labels := range nd.Pods {
			running = append(running, &v1.Pod{ObjectMeta: metav1.ObjectMeta{Name: nd.Name + "-run-" + string(rune('a'+k%26)), Namespace: "default", Labels: labels}})
		}
		ni := framework.NewNodeInfo(running...)
		ni.SetNode(&v1.Node{ObjectMeta: metav1.ObjectMeta{Name: nd.Name}})
		infos[i] = ni
		if !nd.HasScv {
			continue // cache.Get fails for this node: Unschedulable (scheduler.go:81-84)
		}
		obj := &scv.Scv{}
		obj.Name = nd.Name
		obj.Status.CardNumber = uint(nd.CardNumber)
		obj.Status.FreeMemorySum = nd.FreeSum
		obj.Status.TotalMemorySum = nd.TotalSum
		for _, c := range nd.Cards {
			var health string
			if err := json.Unmarshal(c[0], &health); err != nil {
				t.Fatalf("node %s: bad health %s", nd.Name, c[0])
			}
			obj.Status.CardList = append(obj.Status.CardList, scv.Card{Health: health, FreeMemory: u64(t, c[1]), TotalMemory: u64(t, c[2]),
				Clock: uint(u64(t, c[3])), Core: uint(u64(t, c[4])), Bandwidth: uint(u64(t, c[5])), Power: uint(u64(t, c[6]))})
		}
		scvs[i] = obj
		list.Items = append(list.Items, *obj)
	}
	return scvs, list, infos
}

// referenceCase: one pod through the reference, start to finish.
func referenceCase(t *testing.T, labels map[string]string, doc *goldenDoc, scvs []*scv.Scv, list scv.ScvList, infos []*framework.NodeInfo) goldenCase {
	pod := &v1.Pod{ObjectMeta: metav1.ObjectMeta{Name: "pending", Namespace: "default", Labels: labels}}
	out := goldenCase{Labels: labels, Feasible: []string{}, Raw: map[string]int64{}, Norm: map[string]int64{}}
	feasible := make([]bool, len(scvs))
	for i, s := range scvs {
		if s == nil {
			continue
		}
		if ok, number := filter.PodFitsNumber(pod, s); ok { // scheduler.go:85-91
			isFitsMemory, _ := filter.PodFitsMemory(number, pod, s)
			isFitsClock, _ := filter.PodFitsClock(number, pod, s)
			feasible[i] = isFitsMemory && isFitsClock
		}
		if feasible[i] {
			out.Feasible = append(out.Feasible, doc.Nodes[i].Name)
		}
	}
	state := framework.NewCycleState()
	if st := collection.CollectMaxValues(state, pod, list); !st.IsSuccess() { // scheduler.go:102
		t.Fatalf("CollectMaxValues: %s", st.Message())
	}
	d, err := state.Read("Max")
	if err != nil {
		t.Fatalf("CycleState has no Max: %v", err)
	}
	mv := d.(*collection.Data).Value
	out.Max6 = []uint64{uint64(mv.MaxBandwidth), uint64(mv.MaxClock), uint64(mv.MaxCore), mv.MaxFreeMemory, uint64(mv.MaxPower), mv.MaxTotalMemory}
	var scores framework.NodeScoreList
	for i, s := range scvs {
		if !feasible[i] {
			continue
		}
		u, err := score.CalculateScore(s, state, pod, infos[i]) // scheduler.go:124
		if err != nil {
			t.Fatalf("CalculateScore(%s): %v", doc.Nodes[i].Name, err)
		}
		raw := filter.Uint64ToInt64(u) // scheduler.go:128
		out.Raw[doc.Nodes[i].Name] = raw
		scores = append(scores, framework.NodeScore{Name: doc.Nodes[i].Name, Score: raw})
	}
	if len(scores) > 0 {
		if st := (&Yoda{}).NormalizeScore(context.Background(), state, pod, scores); st != nil { // scheduler.go:132-157
			t.Fatalf("NormalizeScore: %s", st.Message())
		}
		best := int64(-1 << 63)
		for _, s := range scores { // list order = node order: the first maximum is the lowest index
			out.Norm[s.Name] = s.Score
			if s.Score > best {
				best = s.Score
				name := s.Name
				out.Chosen = &name
			}
		}
	}
	return out
}

func TestGolden(t *testing.T) {
	if *goldenDir == "" {
		t.Skip("pass -args -golden <repo>/tests/golden")
	}
	for _, fixture := range []string{"kat_cluster.json", "splitmix_cluster.json"} {
		raw, err := ioutil.ReadFile(filepath.Join(*goldenDir, fixture))
		if err != nil {
			t.Fatal(err)
		}
		var doc goldenDoc
		if err := json.Unmarshal(raw, &doc); err != nil {
			t.Fatalf("%s: %v", fixture, err)
		}
		scvs, list, infos := buildCluster(t, &doc)
		ref := goldenDoc{Provenance: "written by go/refgolden/golden_test.go -update from the reference's own packages", Nodes: doc.Nodes}
		for p, want := range doc.Cases {
			got := referenceCase(t, want.Labels, &doc, scvs, list, infos)
			ref.Cases = append(ref.Cases, got)
			if !reflect.DeepEqual(got.Feasible, want.Feasible) {
				t.Errorf("%s pod %d %v: feasible %v, fixture %v", fixture, p, want.Labels, got.Feasible, want.Feasible)
			}
			if !reflect.DeepEqual(got.Max6, want.Max6) {
				t.Errorf("%s pod %d %v: maxima %v, fixture %v", fixture, p, want.Labels, got.Max6, want.Max6)
			}
			if !reflect.DeepEqual(got.Raw, want.Raw) {
				t.Errorf("%s pod %d %v: raw scores differ: %v vs fixture %v", fixture, p, want.Labels, got.Raw, want.Raw)
			}
			if !reflect.DeepEqual(got.Norm, want.Norm) {
				t.Errorf("%s pod %d %v: normalized scores differ: %v vs fixture %v", fixture, p, want.Labels, got.Norm, want.Norm)
			}
			if (got.Chosen == nil) != (want.Chosen == nil) || (got.Chosen != nil && *got.Chosen != *want.Chosen) {
				t.Errorf("%s pod %d %v: chosen node differs", fixture, p, want.Labels)
			}
		}
		for _, kv := range doc.Parser { // filter.StrToUint64 (filter.go:68-74)
			var pair []json.RawMessage
			var s string
			if json.Unmarshal(kv, &pair) != nil || len(pair) != 2 || json.Unmarshal(pair[0], &s) != nil {
				t.Fatalf("bad parser vector %s", kv)
			}
			if got := filter.StrToUint64(s); got != u64(t, pair[1]) {
				t.Errorf("StrToUint64(%q) = %d, fixture %d", s, got, u64(t, pair[1]))
			}
		}
		for _, kv := range doc.U64ToI64 { // filter.Uint64ToInt64 (filter.go:84-86)
			var pair []json.RawMessage
			var want int64
			if json.Unmarshal(kv, &pair) != nil || len(pair) != 2 || json.Unmarshal(pair[1], &want) != nil {
				t.Fatalf("bad uint64_to_int64 vector %s", kv)
			}
			if got := filter.Uint64ToInt64(u64(t, pair[0])); got != want {
				t.Errorf("Uint64ToInt64(%d) = %d, fixture %d", u64(t, pair[0]), got, want)
			}
		}
		for _, kv := range doc.Normalize { // Yoda.NormalizeScore on bare lists
			var pair [][]int64
			if json.Unmarshal(kv, &pair) != nil || len(pair) != 2 {
				t.Fatalf("bad normalize vector %s", kv)
			}
			var scores framework.NodeScoreList
			for _, v := range pair[0] {
				scores = append(scores, framework.NodeScore{Name: "n", Score: v})
			}
			(&Yoda{}).NormalizeScore(context.Background(), nil, &v1.Pod{}, scores)
			for k, s := range scores {
				if s.Score != pair[1][k] {
					t.Errorf("NormalizeScore(%v)[%d] = %d, fixture %d", pair[0], k, s.Score, pair[1][k])
				}
			}
		}
		if *update {
			b, err := json.Marshal(&ref)
			if err != nil {
				t.Fatal(err)
			}
			if err := ioutil.WriteFile(filepath.Join(*goldenDir, "ref_"+fixture), b, 0644); err != nil {
				t.Fatal(err)
			}
		}
	}
}
