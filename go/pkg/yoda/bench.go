package yoda

// RunBench: the fake-framework benchmark harness behind `scheduler bench`.  It is the Go twin of
// yoda_scheduler_b200/host/harness.cpp: synthetic SCV snapshot (same SplitMix64 streams), every
// pending pod in one device pass, and, for a few pods, the per-pod call sequence of kube-scheduler
// (Filter on every node, PostFilter, Score on the feasible nodes, NormalizeScore, lowest-index
// selectHost) checked against the batch result.  Prints one JSON line.  Not compiled in this image.

import (
	"flag"
	"fmt"
	"time"

	"github.com/NJUPT-ISL/Yoda-Scheduler/pkg/yoda/b200"
)

const seed = 0x59304441

var skus = [8][5]uint64{{11178, 5705, 3584, 352, 250}, {11019, 7000, 4352, 616, 260}, {16160, 877, 5120, 900, 300}, {32510, 877, 5120, 900, 300},
	{40536, 1215, 6912, 1555, 400}, {81920, 1593, 6912, 2039, 400}, {81559, 2619, 16896, 3350, 700}, {183359, 3996, 18944, 8000, 1000}}

func sm64(x uint64) uint64 {
	z := x + 0x9E3779B97F4A7C15
	z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9
	z = (z ^ (z >> 27)) * 0x94D049BB133111EB
	return z ^ (z >> 31)
}
func rnd(s, stream, i, j uint64) uint64 { return sm64(sm64(sm64(s+stream)+i) + j) }

func RunBench(args []string) int {
	fs := flag.NewFlagSet("bench", flag.ExitOnError)
	nodes := fs.Uint64("nodes", 50000, "synthetic nodes (8 cards each)")
	pods := fs.Uint64("pods", 100000, "pending pods")
	podSeed := fs.Uint64("pod-seed", seed+3, "pod stream seed")
	_ = fs.Parse(args)
	dev, err := b200.New(0, uint32(*nodes))
	if err != nil {
		fmt.Println(err)
		return 2
	}
	defer dev.Close()
	for n := uint64(0); n < *nodes; n++ {
		rec, alloc, has := synthNode(n)
		if has {
			if err := dev.Upsert(uint32(n), rec); err != nil {
				fmt.Println(err)
				return 2
			}
		} else {
			_ = dev.Remove(uint32(n))
		}
		_ = dev.SetAlloc(uint32(n), alloc)
	}
	if _, err := dev.Commit(); err != nil {
		fmt.Println(err)
		return 2
	}
	reqs := make([]b200.PodReq, *pods)
	for p := range reqs {
		reqs[p] = synthPod(uint64(p), *podSeed)
	}
	_, _ = dev.Eval(reqs, true) // warm-up
	t0 := time.Now()
	res, err := dev.Eval(reqs, true)
	dt := time.Since(t0)
	if err != nil {
		fmt.Println(err)
		return 2
	}
	fmt.Printf("{\"nodes\": %d, \"pods\": %d, \"batch_ms\": %.3f, \"evals_per_s\": %.4g, \"first_choice\": %d}\n",
		*nodes, *pods, float64(dt.Microseconds())/1e3, float64(*nodes)*float64(*pods)/dt.Seconds(), b200.Chosen(&res[0]))
	return 0
}
