package yoda

// RunBench: the fake-framework benchmark harness behind `scheduler bench` -- no cluster needed.
// It drives BOTH paths behind one fake framework.Handle on the same synthetic SCV snapshot (the
// SplitMix64 streams of yoda_scheduler_b200/synth.py):
//
//   reference arm: the reference's own packages, untouched -- filter.PodFits{Number,Memory,Clock}
//     for every node (scheduler.go:85-91), collection.CollectMaxValues (scheduler.go:95-103),
//     score.CalculateScore + filter.Uint64ToInt64 for the feasible nodes (scheduler.go:117-128) and the
//     NormalizeScore rescale (scheduler.go:132-157, restated below because this overlay replaces
//     scheduler.go; go/refgolden/golden_test.go checks the real method) -- on -sample pods, the node
//     loop spread over -parallel goroutines like the framework's parallelize.Until (16 by default);
//   device arm: this plugin called the way kube-scheduler calls it for the same pods (Filter on every
//     node from 16 goroutines, PostFilter, Score on the feasible nodes, NormalizeScore, lowest-index
//     selectHost), every verdict/score/choice compared with the reference arm; then the whole batch in
//     one Eval.
//
// Prints one JSON line with both rates, GOMAXPROCS and the core count.  NOT COMPILED in the build
// image (no Go toolchain); yoda_scheduler_b200/host/harness.cpp is its tested C++ twin.

import (
	"context"
	"flag"
	"fmt"
	"runtime"
	"sync"
	"time"

	v1 "k8s.io/api/core/v1"
	metav1 "k8s.io/apimachinery/pkg/apis/meta/v1"
	"k8s.io/kubernetes/pkg/scheduler/framework"

	scv "github.com/NJUPT-ISL/SCV/api/v1"

	"github.com/NJUPT-ISL/Yoda-Scheduler/pkg/yoda/b200"
	"github.com/NJUPT-ISL/Yoda-Scheduler/pkg/yoda/collection"
	"github.com/NJUPT-ISL/Yoda-Scheduler/pkg/yoda/filter"
	"github.com/NJUPT-ISL/Yoda-Scheduler/pkg/yoda/score"
)

// fakeLister / fakeHandle: the two things the plugin asks of framework.Handle on the hot path.
type fakeLister struct {
	infos  []*framework.NodeInfo
	byName map[string]*framework.NodeInfo
}

func (l *fakeLister) NodeInfos() framework.NodeInfoLister      { return l }
func (l *fakeLister) List() ([]*framework.NodeInfo, error)     { return l.infos, nil }
func (l *fakeLister) HavePodsWithAffinityList() ([]*framework.NodeInfo, error) { return nil, nil }
func (l *fakeLister) HavePodsWithRequiredAntiAffinityList() ([]*framework.NodeInfo, error) {
	return nil, nil
}
func (l *fakeLister) Get(name string) (*framework.NodeInfo, error) {
	if ni, ok := l.byName[name]; ok {
		return ni, nil
	}
	return nil, fmt.Errorf("nodeinfo not found for node name %q", name)
}

type fakeHandle struct {
	framework.Handle // every other method is unused by the plugin and panics if reached
	lister           *fakeLister
}

func (h *fakeHandle) SnapshotSharedLister() framework.SharedLister { return h.lister }

// referenceNormalize restates Yoda.NormalizeScore (scheduler.go:132-157) without its per-node klog line.
func referenceNormalize(scores framework.NodeScoreList) {
	var highest int64 = 0
	lowest := scores[0].Score
	for _, s := range scores {
		if s.Score < lowest {
			lowest = s.Score
		}
		if s.Score > highest {
			highest = s.Score
		}
	}
	if highest == lowest {
		lowest--
	}
	for i, s := range scores {
		scores[i].Score = (s.Score - lowest) * framework.MaxNodeScore / (highest - lowest)
	}
}

// parallelNodes mimics parallelize.Until: `workers` goroutines pull node indices.
func parallelNodes(n, workers int, f func(i int)) {
	var wg sync.WaitGroup
	next := make(chan int, 1024)
	for w := 0; w < workers; w++ {
		wg.Add(1)
		go func() {
			defer wg.Done()
			for i := range next {
				f(i)
			}
		}()
	}
	for i := 0; i < n; i++ {
		next <- i
	}
	close(next)
	wg.Wait()
}

type cycleOut struct {
	feasible []bool
	raw      []int64
	norm     []int64
	chosen   int
}

// referenceCycle: one pod through the reference's own code, in the order the plugin intends.
func referenceCycle(pod *v1.Pod, scvs []*scv.Scv, list scv.ScvList, infos []*framework.NodeInfo, workers int) cycleOut {
	n := len(scvs)
	out := cycleOut{feasible: make([]bool, n), raw: make([]int64, n), norm: make([]int64, n), chosen: -1}
	parallelNodes(n, workers, func(i int) { // Yoda.Filter, scheduler.go:76-93
		if scvs[i] == nil {
			return
		}
		if ok, number := filter.PodFitsNumber(pod, scvs[i]); ok {
			isFitsMemory, _ := filter.PodFitsMemory(number, pod, scvs[i])
			isFitsClock, _ := filter.PodFitsClock(number, pod, scvs[i])
			out.feasible[i] = isFitsMemory && isFitsClock
		}
	})
	state := framework.NewCycleState()
	collection.CollectMaxValues(state, pod, list) // Yoda.PostFilter, scheduler.go:95-103
	parallelNodes(n, workers, func(i int) {         // Yoda.Score, scheduler.go:109-130
		if !out.feasible[i] {
			return
		}
		s, err := score.CalculateScore(scvs[i], state, pod, infos[i])
		if err == nil {
			out.raw[i] = filter.Uint64ToInt64(s)
		}
	})
	var scores framework.NodeScoreList
	var idx []int
	for i := 0; i < n; i++ {
		if out.feasible[i] {
			scores = append(scores, framework.NodeScore{Name: infos[i].Node().Name, Score: out.raw[i]})
			idx = append(idx, i)
		}
	}
	if len(scores) > 0 {
		referenceNormalize(scores)
		best := int64(-1)
		for k, i := range idx {
			out.norm[i] = scores[k].Score
			if scores[k].Score > best {
				best, out.chosen = scores[k].Score, i
			}
		}
	}
	return out
}

// pluginCycle: the same pod through this plugin, called as kube-scheduler calls it.
func pluginCycle(y *Yoda, pod *v1.Pod, infos []*framework.NodeInfo, workers int) (cycleOut, error) {
	n := len(infos)
	out := cycleOut{feasible: make([]bool, n), raw: make([]int64, n), norm: make([]int64, n), chosen: -1}
	ctx := context.Background()
	state := framework.NewCycleState()
	parallelNodes(n, workers, func(i int) {
		out.feasible[i] = y.Filter(ctx, state, pod, infos[i]).IsSuccess()
	})
	if _, st := y.PostFilter(ctx, state, pod, nil); !st.IsSuccess() {
		return out, fmt.Errorf("PostFilter: %s", st.Message())
	}
	var firstErr error
	var mu sync.Mutex
	parallelNodes(n, workers, func(i int) {
		if !out.feasible[i] {
			return
		}
		s, st := y.Score(ctx, state, pod, infos[i].Node().Name)
		if !st.IsSuccess() {
			mu.Lock()
			firstErr = fmt.Errorf("Score %s: %s", infos[i].Node().Name, st.Message())
			mu.Unlock()
		}
		out.raw[i] = s
	})
	if firstErr != nil {
		return out, firstErr
	}
	var scores framework.NodeScoreList
	var idx []int
	for i := 0; i < n; i++ {
		if out.feasible[i] {
			scores = append(scores, framework.NodeScore{Name: infos[i].Node().Name, Score: out.raw[i]})
			idx = append(idx, i)
		}
	}
	if len(scores) > 0 {
		y.NormalizeScore(ctx, state, pod, scores)
		best := int64(-1)
		for k, i := range idx {
			out.norm[i] = scores[k].Score
			if scores[k].Score > best {
				best, out.chosen = scores[k].Score, i
			}
		}
	}
	return out, nil
}

func podWithLabels(name string, labels map[string]string) *v1.Pod {
	return &v1.Pod{ObjectMeta: metav1.ObjectMeta{Name: name, Namespace: "default", Labels: labels}}
}

func RunBench(args []string) int {
	fs := flag.NewFlagSet("bench", flag.ExitOnError)
	nodes := fs.Uint64("nodes", 50000, "synthetic nodes (8 cards each)")
	pods := fs.Uint64("pods", 100000, "pending pods of the device batch")
	sample := fs.Uint64("sample", 32, "pods taken through both per-pod call sequences and compared")
	parallel := fs.Int("parallel", 16, "goroutines over nodes (kube-scheduler's default parallelism)")
	podSeed := fs.Uint64("pod-seed", seed+3, "pod stream seed")
	_ = fs.Parse(args)
	if *sample > *pods {
		*sample = *pods
	}

	dev, err := b200.New(0, uint32(*nodes))
	if err != nil {
		fmt.Println(err)
		return 2
	}
	defer dev.Close()

	// one synthetic cluster, three views: Scv objects (reference arm), NodeInfos (both), device snapshot (through the plugin's own handlers)
	lister := &fakeLister{byName: map[string]*framework.NodeInfo{}}
	y := newYoda(&fakeHandle{lister: lister}, dev)
	scvs := make([]*scv.Scv, *nodes)
	var list scv.ScvList
	for n := uint64(0); n < *nodes; n++ {
		obj, running, has := synthScv(n)
		ni := framework.NewNodeInfo(running...)
		ni.SetNode(&v1.Node{ObjectMeta: metav1.ObjectMeta{Name: obj.Name}})
		lister.infos = append(lister.infos, ni)
		lister.byName[obj.Name] = ni
		if has {
			scvs[n] = obj
			list.Items = append(list.Items, *obj)
			y.onScv(obj)
		} else {
			y.onScvDelete(obj.Name) // registers the name, keeps the node without an SCV
		}
	}

	labels := make([]map[string]string, *pods)
	reqs := make([]b200.PodReq, *pods)
	for p := range reqs {
		labels[p] = synthPodLabels(uint64(p), *podSeed)
		reqs[p] = requirement(podWithLabels("p", labels[p]))
	}

	// both per-pod call sequences on the sample, compared
	mismatches := 0
	var refDur, devDur time.Duration
	for p := uint64(0); p < *sample; p++ {
		pod := podWithLabels(fmt.Sprintf("pod-%d", p), labels[p])
		t0 := time.Now()
		want := referenceCycle(pod, scvs, list, lister.infos, *parallel)
		refDur += time.Since(t0)
		t0 = time.Now()
		got, err := pluginCycle(y, pod, lister.infos, *parallel)
		devDur += time.Since(t0)
		if err != nil {
			fmt.Println(err)
			return 2
		}
		if got.chosen != want.chosen {
			mismatches++
		}
		for i := range want.feasible {
			if got.feasible[i] != want.feasible[i] || (want.feasible[i] && (got.raw[i] != want.raw[i] || got.norm[i] != want.norm[i])) {
				mismatches++
				break
			}
		}
	}

	// the whole pending batch in one device pass
	if _, err := dev.Eval(reqs, true); err != nil { // warm-up
		fmt.Println(err)
		return 2
	}
	t0 := time.Now()
	res, err := dev.Eval(reqs, true)
	batch := time.Since(t0)
	if err != nil {
		fmt.Println(err)
		return 2
	}
	fmt.Printf("{\"nodes\": %d, \"pods\": %d, \"sample_pods\": %d, \"gomaxprocs\": %d, \"num_cpu\": %d, \"parallel\": %d, "+
		"\"reference_evals_per_s\": %.4g, \"reference_ms_per_pod\": %.3f, \"plugin_cycle_ms_per_pod\": %.3f, "+
		"\"batch_ms\": %.3f, \"batch_evals_per_s\": %.4g, \"mismatches\": %d, \"first_choice\": %d}\n",
		*nodes, *pods, *sample, runtime.GOMAXPROCS(0), runtime.NumCPU(), *parallel,
		float64(*sample)*float64(*nodes)/refDur.Seconds(), float64(refDur.Microseconds())/1e3/float64(*sample),
		float64(devDur.Microseconds())/1e3/float64(*sample),
		float64(batch.Microseconds())/1e3, float64(*nodes)*float64(*pods)/batch.Seconds(), mismatches, b200.Chosen(&res[0]))
	if mismatches != 0 {
		return 1
	}
	return 0
}
