// Package yoda: the plugin surface kube-scheduler sees (name "yoda"; QueueSort, Filter,
// PostFilter, Score, ScoreExtensions), with the per-(pod,node) arithmetic moved to the device.
//
// This file replaces pkg/yoda/scheduler.go of the reference checkout (see ../../apply_overlay.sh);
// pkg/yoda/filter, collection, score and sort stay the reference's.  The controller-runtime cache
// is still what watches Scv objects, but instead of being read (and deep-copied) on every call it
// feeds the device-resident snapshot through event handlers.
//
// NOT COMPILED in the build image (no Go toolchain there); the same logic is tested through the
// C ABI by the C++ mirror yoda_scheduler_b200/host/yoda_plugin.cpp.
package yoda

import (
	"context"
	"fmt"
	"strconv"
	"sync"
	"unsafe"

	v1 "k8s.io/api/core/v1"
	"k8s.io/apimachinery/pkg/runtime"
	toolscache "k8s.io/client-go/tools/cache"
	"k8s.io/klog"
	"k8s.io/kubernetes/pkg/scheduler/framework"
	ctrl "sigs.k8s.io/controller-runtime"

	scv "github.com/NJUPT-ISL/SCV/api/v1"

	"github.com/NJUPT-ISL/Yoda-Scheduler/pkg/yoda/b200"
	"github.com/NJUPT-ISL/Yoda-Scheduler/pkg/yoda/collection"
	"github.com/NJUPT-ISL/Yoda-Scheduler/pkg/yoda/sort"
)

const Name = "yoda"

const memoKey = "yoda/b200"

var (
	_ framework.QueueSortPlugin  = &Yoda{}
	_ framework.FilterPlugin     = &Yoda{}
	_ framework.PostFilterPlugin = &Yoda{}
	_ framework.ScorePlugin      = &Yoda{}
	_ framework.ScoreExtensions  = &Yoda{}

	scheme = runtime.NewScheme()
)

// Yoda keeps the node-name <-> device-index dictionary and the host mirror of what the device
// holds per node.  mu guards every field below it: informer goroutines and evaluations take it for
// writing.  The per-node calls of a cycle (Filter / Score from the 16 framework goroutines) do not
// take it: they read the cycle's memo, which pins the dictionary as it was at the evaluation.  The
// dictionary is copied before a new name is added while a memo may still read it (indexPinned).
type Yoda struct {
	handle framework.Handle
	dev    *b200.Ctx

	mu          sync.RWMutex
	index       map[string]uint32 // node name -> dense device index
	indexPinned bool              // some cycle's memo holds `index`: copy before the next insert
	names   []string
	hasScv  []bool
	reject  []string // why the node's last Scv was refused by the device ("" = accepted)
	alloc   []uint64 // sum of StrToUint64(scv/memory) over NodeInfo.Pods, as last sent to the device
	gen     []int64  // NodeInfo.Generation the alloc sum was computed for
	genSeen []bool
	dirty   bool
}

// memo is what one cycle keeps in CycleState after its single device pass: a frozen view, so the
// later Filter/Score/NormalizeScore calls of the cycle never look at state an informer may be changing.
type memo struct {
	res    b200.PodResult
	n      uint32            // nodes the evaluation covered; indices >= n were registered later
	index  map[string]uint32 // the dictionary at evaluation time; never written while a memo holds it
	has    []bool   // copy of hasScv[:n] at evaluation time
	reject []string // copy of reject[:n]
	bits   []uint32
	raw    []int64
	norm   []uint8
}

func (m *memo) Clone() framework.StateData { return m }
func (m *memo) known(i uint32) bool        { return i < m.n && m.has[i] }
func (m *memo) feasible(i uint32) bool     { return i < m.n && (m.bits[i>>5]>>(i&31))&1 == 1 }

func (y *Yoda) Name() string { return Name }

// atoiOrZero is the label rule of the reference (filter.go:60-74): strconv.Atoi, any error -> 0,
// then the uint conversion.
func atoiOrZero(s string) uint64 {
	i, err := strconv.Atoi(s)
	if err != nil {
		return 0
	}
	return uint64(i)
}

func requirement(pod *v1.Pod) b200.PodReq {
	l := pod.GetLabels()
	num, hasNum := l["scv/number"]
	mem, hasMem := l["scv/memory"]
	clk, hasClk := l["scv/clock"]
	return b200.MakeReq(hasNum, hasMem, hasClk, atoiOrZero(num), atoiOrZero(mem), atoiOrZero(clk))
}

// newYoda builds the plugin around an existing device context (New, and the bench harness).
func newYoda(h framework.Handle, dev *b200.Ctx) *Yoda {
	return &Yoda{handle: h, dev: dev, index: map[string]uint32{}}
}

func New(_ runtime.Object, h framework.Handle) (framework.Plugin, error) {
	if err := scv.AddToScheme(scheme); err != nil {
		return nil, err
	}
	dev, err := b200.New(0, 1<<20)
	if err != nil {
		return nil, err
	}
	y := newYoda(h, dev)

	mgr, err := ctrl.NewManager(ctrl.GetConfigOrDie(), ctrl.Options{Scheme: scheme, MetricsBindAddress: "", LeaderElection: false, Port: 9443})
	if err != nil {
		return nil, err
	}
	inf, err := mgr.GetCache().GetInformer(context.Background(), &scv.Scv{})
	if err != nil {
		return nil, err
	}
	inf.AddEventHandler(toolscache.ResourceEventHandlerFuncs{
		AddFunc:    func(o interface{}) { y.onScv(o.(*scv.Scv)) },
		UpdateFunc: func(_, o interface{}) { y.onScv(o.(*scv.Scv)) },
		DeleteFunc: func(o interface{}) {
			if t, ok := o.(toolscache.DeletedFinalStateUnknown); ok { // tombstone of a missed delete
				o = t.Obj
			}
			if s, ok := o.(*scv.Scv); ok {
				y.onScvDelete(s.Name)
			}
		},
	})
	// CalculateAllocateScore reads NodeInfo.Pods of the scheduler's snapshot (algorithm.go:74-80), which
	// holds assumed pods as well as bound ones.  No Pod informer can reproduce that, so the sums are taken
	// from the very same snapshot at the start of each cycle (syncAllocLocked), node generation by generation.
	go func() {
		if err := mgr.Start(ctrl.SetupSignalHandler()); err != nil {
			klog.Error(err)
			panic(err)
		}
	}()
	return y, nil
}

// indexOfLocked: callers hold mu for writing.
func (y *Yoda) indexOfLocked(name string) uint32 {
	if i, ok := y.index[name]; ok {
		return i
	}
	i := uint32(len(y.names))
	if y.indexPinned { // a cycle still looks names up in the current map: leave it untouched
		next := make(map[string]uint32, len(y.index)+1)
		for k, v := range y.index {
			next[k] = v
		}
		y.index, y.indexPinned = next, false
	}
	y.index[name] = i
	y.names = append(y.names, name)
	y.hasScv = append(y.hasScv, false)
	y.reject = append(y.reject, "")
	y.alloc = append(y.alloc, 0)
	y.gen = append(y.gen, 0)
	y.genSeen = append(y.genSeen, false)
	return i
}

func (y *Yoda) onScv(s *scv.Scv) {
	y.mu.Lock()
	defer y.mu.Unlock()
	i := y.indexOfLocked(s.Name)
	rec := b200.RecordFromScv(s) // healthy bit g = (CardList[g].Health == "Healthy")
	if err := y.dev.Upsert(i, rec); err != nil {
		// more than 8 cards, or TotalMemorySum == 0 (where the reference would panic, algorithm.go:71,86):
		// the node must not keep answering from its previous record
		klog.Errorf("Upsert SCV Error: %v", err)
		_ = y.dev.Remove(i)
		y.hasScv[i], y.reject[i], y.dirty = false, err.Error(), true
		return
	}
	y.hasScv[i], y.reject[i], y.dirty = true, "", true
}

func (y *Yoda) onScvDelete(name string) {
	y.mu.Lock()
	defer y.mu.Unlock()
	i := y.indexOfLocked(name)
	_ = y.dev.Remove(i)
	y.hasScv[i], y.reject[i], y.dirty = false, "", true
}

// syncAllocLocked brings the per-node allocateMemorySum in line with the scheduler's snapshot: the sum of
// StrToUint64(scv/memory) over NodeInfo.Pods (algorithm.go:75-80).  Only nodes whose NodeInfo.Generation
// moved since the last cycle are re-summed.  Callers hold mu for writing.
func (y *Yoda) syncAllocLocked() {
	if y.handle == nil {
		return
	}
	infos, err := y.handle.SnapshotSharedLister().NodeInfos().List()
	if err != nil {
		klog.Errorf("List NodeInfos Error: %v", err)
		return
	}
	for _, ni := range infos {
		node := ni.Node()
		if node == nil {
			continue
		}
		i := y.indexOfLocked(node.Name)
		if y.genSeen[i] && y.gen[i] == ni.Generation {
			continue
		}
		var sum uint64
		for _, pi := range ni.Pods {
			if mem, ok := pi.Pod.GetLabels()["scv/memory"]; ok {
				sum += atoiOrZero(mem)
			}
		}
		y.gen[i], y.genSeen[i] = ni.Generation, true
		if sum != y.alloc[i] {
			y.alloc[i] = sum
			if err := y.dev.SetAlloc(i, sum); err != nil {
				klog.Errorf("SetAlloc Error: %v", err)
			}
			y.dirty = true
		}
	}
}

// cycle returns the memo of this scheduling cycle, evaluating the pod once on first use.
func (y *Yoda) cycle(state *framework.CycleState, pod *v1.Pod) (*memo, error) {
	state.RLock()
	d, err := state.Read(memoKey)
	state.RUnlock()
	if err == nil {
		return d.(*memo), nil
	}
	state.Lock()
	defer state.Unlock()
	if d, err := state.Read(memoKey); err == nil {
		return d.(*memo), nil
	}
	y.mu.Lock()
	defer y.mu.Unlock()
	y.syncAllocLocked()
	if y.dirty {
		if _, err := y.dev.Commit(); err != nil {
			return nil, err
		}
		y.dirty = false
	}
	res, err := y.dev.Eval([]b200.PodReq{requirement(pod)}, true)
	if err != nil {
		return nil, err
	}
	n := y.dev.NodeCount()
	if int(n) > len(y.hasScv) {
		n = uint32(len(y.hasScv))
	}
	y.indexPinned = true
	m := &memo{res: res[0], n: n, index: y.index, bits: make([]uint32, (n+31)/32+1), raw: make([]int64, n+1), norm: make([]uint8, n+1)}
	m.has = append([]bool(nil), y.hasScv[:n]...)
	m.reject = append([]string(nil), y.reject[:n]...)
	if n > 0 {
		if err = y.dev.Row(0, b200.RowFeasible, unsafe.Pointer(&m.bits[0]), int((n+31)/32)*4); err == nil {
			err = y.dev.Row(0, b200.RowRaw, unsafe.Pointer(&m.raw[0]), int(n)*8)
		}
		if err == nil {
			err = y.dev.Row(0, b200.RowNorm, unsafe.Pointer(&m.norm[0]), int(n))
		}
		if err != nil {
			return nil, err
		}
	}
	state.Write(memoKey, m)
	return m, nil
}

func notFound(name string) string { return "Scv.core.run-linux.com \"" + name + "\" not found" }

func (y *Yoda) Filter(_ context.Context, state *framework.CycleState, pod *v1.Pod, node *framework.NodeInfo) *framework.Status {
	name := node.Node().Name
	m, err := y.cycle(state, pod)
	if err != nil {
		return framework.NewStatus(framework.Error, err.Error())
	}
	i, ok := m.index[name]
	if !ok || !m.known(i) {
		why := notFound(name)
		if ok && i < m.n && m.reject[i] != "" {
			why = "Scv rejected: " + m.reject[i]
		}
		return framework.NewStatus(framework.Unschedulable, "Node:"+name+" "+why)
	}
	if m.feasible(i) {
		return framework.NewStatus(framework.Success, "")
	}
	return framework.NewStatus(framework.Unschedulable, "Node:"+name)
}

func (y *Yoda) PostFilter(_ context.Context, state *framework.CycleState, pod *v1.Pod, _ framework.NodeToStatusMap) (*framework.PostFilterResult, *framework.Status) {
	m, err := y.cycle(state, pod)
	if err != nil {
		return &framework.PostFilterResult{}, framework.NewStatus(framework.Error, err.Error())
	}
	mx := b200.Maxima(&m.res) // order of collection.MaxValue: bandwidth, clock, core, free, power, total
	data := collection.Data{Value: collection.MaxValue{
		MaxBandwidth: uint(mx[0]), MaxClock: uint(mx[1]), MaxCore: uint(mx[2]),
		MaxFreeMemory: mx[3], MaxPower: uint(mx[4]), MaxTotalMemory: mx[5],
	}}
	state.Lock()
	state.Write("Max", &data)
	state.Unlock()
	return &framework.PostFilterResult{}, framework.NewStatus(framework.Success, "")
}

func (y *Yoda) Less(a, b *framework.QueuedPodInfo) bool { return sort.Less(a, b) }

func (y *Yoda) Score(_ context.Context, state *framework.CycleState, p *v1.Pod, nodeName string) (int64, *framework.Status) {
	if _, err := y.handle.SnapshotSharedLister().NodeInfos().Get(nodeName); err != nil {
		return 0, framework.NewStatus(framework.Error, fmt.Sprintf("getting node %q from Snapshot: %v", nodeName, err))
	}
	m, err := y.cycle(state, p)
	if err != nil {
		return 0, framework.NewStatus(framework.Error, fmt.Sprintf("Score Node Error: %v", err))
	}
	i, ok := m.index[nodeName]
	if !ok || !m.known(i) { // scheduler.go:117-122; also a node registered after this cycle's evaluation
		return 0, framework.NewStatus(framework.Error, "Score Node Error: "+notFound(nodeName))
	}
	state.RLock()
	_, err = state.Read("Max") // algorithm.go:29-33
	state.RUnlock()
	if err != nil {
		return 0, framework.NewStatus(framework.Error, fmt.Sprintf("Score Node Error: %v", err))
	}
	return m.raw[i], framework.NewStatus(framework.Success, "")
}

// NormalizeScore: when the list is Yoda's own feasible set with Yoda's scores the device already
// normalized it; when other plugins removed nodes, rescale the given list on the device.
func (y *Yoda) NormalizeScore(_ context.Context, state *framework.CycleState, _ *v1.Pod, scores framework.NodeScoreList) *framework.Status {
	_ = scores[0] // an empty list panics, as upstream (scheduler.go:135)
	var m *memo
	state.RLock()
	if d, err := state.Read(memoKey); err == nil {
		m = d.(*memo)
	}
	state.RUnlock()
	own := m != nil && uint32(len(scores)) == b200.Feasible(&m.res)
	idx := make([]uint32, len(scores))
	if own {
		for k := 0; own && k < len(scores); k++ {
			i, ok := m.index[scores[k].Name]
			// norm == 255 marks a rescale that wrapped outside [0,100]: the reference returns the wrapped int64,
			// which only the exact list path below reproduces
			own = ok && m.feasible(i) && m.raw[i] == scores[k].Score && m.norm[i] != 255
			idx[k] = i
		}
	}
	if own {
		for k := range scores {
			scores[k].Score = int64(m.norm[idx[k]])
		}
		return nil
	}
	vals := make([]int64, len(scores))
	for k := range scores {
		vals[k] = scores[k].Score
	}
	y.mu.Lock()
	err := y.dev.NormalizeScores(vals)
	y.mu.Unlock()
	if err != nil {
		return framework.NewStatus(framework.Error, err.Error())
	}
	for k := range scores {
		scores[k].Score = vals[k]
	}
	return nil
}

func (y *Yoda) ScoreExtensions() framework.ScoreExtensions { return y }
