// Package yoda: the plugin surface kube-scheduler sees (name "yoda"; QueueSort, Filter,
// PostFilter, Score, ScoreExtensions), with the per-(pod,node) arithmetic moved to the device.
// The controller-runtime cache is still what watches Scv objects, but instead of being read on
// every call it feeds the device-resident snapshot through event handlers.
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

type Yoda struct {
	handle framework.Handle
	dev    *b200.Ctx

	mu     sync.Mutex
	index  map[string]uint32 // node name -> dense device index
	names  []string
	hasScv []bool
	alloc  []uint64 // sum of StrToUint64(scv/memory) over the node's pods
	dirty  bool
}

// memo is what one cycle keeps in CycleState after its single device pass.
type memo struct {
	res  b200.PodResult
	n    uint32
	bits []uint32
	raw  []int64
	norm []uint8
}

func (m *memo) Clone() framework.StateData { return m }
func (m *memo) feasible(i uint32) bool     { return i < m.n && (m.bits[i>>5]>>(i&31))&1 == 1 }

func (y *Yoda) Name() string { return Name }

// atoiOrZero is the label rule of the reference: strconv.Atoi, any error -> 0, then uint conversion.
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

func New(_ runtime.Object, h framework.Handle) (framework.Plugin, error) {
	if err := scv.AddToScheme(scheme); err != nil {
		return nil, err
	}
	dev, err := b200.New(0, 1<<20)
	if err != nil {
		return nil, err
	}
	y := &Yoda{handle: h, dev: dev, index: map[string]uint32{}}

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
			if s, ok := o.(*scv.Scv); ok {
				y.onScvDelete(s.Name)
			}
		},
	})
	// running pods change CalculateAllocateScore: follow them through the scheduler's own pod informer
	h.SharedInformerFactory().Core().V1().Pods().Informer().AddEventHandler(toolscache.ResourceEventHandlerFuncs{
		AddFunc:    func(o interface{}) { y.onPod(o.(*v1.Pod), +1) },
		DeleteFunc: func(o interface{}) {
			if p, ok := o.(*v1.Pod); ok {
				y.onPod(p, -1)
			}
		},
	})
	go func() {
		if err := mgr.Start(ctrl.SetupSignalHandler()); err != nil {
			klog.Error(err)
			panic(err)
		}
	}()
	return y, nil
}

func (y *Yoda) indexOf(name string) uint32 {
	if i, ok := y.index[name]; ok {
		return i
	}
	i := uint32(len(y.names))
	y.index[name] = i
	y.names = append(y.names, name)
	y.hasScv = append(y.hasScv, false)
	y.alloc = append(y.alloc, 0)
	return i
}

func (y *Yoda) onScv(s *scv.Scv) {
	y.mu.Lock()
	defer y.mu.Unlock()
	i := y.indexOf(s.Name)
	rec := b200.RecordFromScv(s) // healthy bit g = (CardList[g].Health == "Healthy")
	if err := y.dev.Upsert(i, rec); err != nil {
		klog.Errorf("Upsert SCV Error: %v", err)
		return
	}
	y.hasScv[i], y.dirty = true, true
}

func (y *Yoda) onScvDelete(name string) {
	y.mu.Lock()
	defer y.mu.Unlock()
	i := y.indexOf(name)
	_ = y.dev.Remove(i)
	y.hasScv[i], y.dirty = false, true
}

func (y *Yoda) onPod(p *v1.Pod, sign int) {
	mem, ok := p.GetLabels()["scv/memory"]
	if !ok || p.Spec.NodeName == "" {
		return
	}
	y.mu.Lock()
	defer y.mu.Unlock()
	i := y.indexOf(p.Spec.NodeName)
	if sign > 0 {
		y.alloc[i] += atoiOrZero(mem)
	} else {
		y.alloc[i] -= atoiOrZero(mem)
	}
	_ = y.dev.SetAlloc(i, y.alloc[i])
	y.dirty = true
}

// cycle returns the memo of this scheduling cycle, evaluating the pod once on first use.
func (y *Yoda) cycle(state *framework.CycleState, pod *v1.Pod) (*memo, error) {
	state.Lock()
	defer state.Unlock()
	if d, err := state.Read(memoKey); err == nil {
		return d.(*memo), nil
	}
	y.mu.Lock()
	defer y.mu.Unlock()
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
	m := &memo{res: res[0], n: n, bits: make([]uint32, (n+31)/32+1), raw: make([]int64, n+1), norm: make([]uint8, n+1)}
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

func (y *Yoda) Filter(_ context.Context, state *framework.CycleState, pod *v1.Pod, node *framework.NodeInfo) *framework.Status {
	name := node.Node().Name
	m, err := y.cycle(state, pod)
	if err != nil {
		return framework.NewStatus(framework.Error, err.Error())
	}
	i, ok := y.index[name]
	if !ok || !y.hasScv[i] {
		return framework.NewStatus(framework.Unschedulable, "Node:"+name+" Scv.core.run-linux.com \""+name+"\" not found")
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
	i, ok := y.index[nodeName]
	if !ok || !y.hasScv[i] {
		return 0, framework.NewStatus(framework.Error, fmt.Sprintf("Score Node Error: Scv.core.run-linux.com %q not found", nodeName))
	}
	if _, err := state.Read("Max"); err != nil {
		return 0, framework.NewStatus(framework.Error, fmt.Sprintf("Score Node Error: %v", err))
	}
	m, err := y.cycle(state, p)
	if err != nil {
		return 0, framework.NewStatus(framework.Error, fmt.Sprintf("Score Node Error: %v", err))
	}
	return m.raw[i], framework.NewStatus(framework.Success, "")
}

// NormalizeScore: when the list is Yoda's own feasible set with Yoda's scores the device already
// normalized it; when other plugins removed nodes, rescale the given list on the device.
func (y *Yoda) NormalizeScore(_ context.Context, state *framework.CycleState, _ *v1.Pod, scores framework.NodeScoreList) *framework.Status {
	_ = scores[0] // an empty list panics, as upstream
	var m *memo
	if d, err := state.Read(memoKey); err == nil {
		m = d.(*memo)
	}
	own := m != nil && uint32(len(scores)) == b200.Feasible(&m.res)
	for k := 0; own && k < len(scores); k++ {
		i, ok := y.index[scores[k].Name]
		own = ok && m.feasible(i) && m.raw[i] == scores[k].Score
	}
	if own {
		for k := range scores {
			scores[k].Score = int64(m.norm[y.index[scores[k].Name]])
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
