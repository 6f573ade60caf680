package yoda

// Synthetic cluster and pending pods of the bench harness: Go twins of yoda_scheduler_b200/synth.py
// (counter-based SplitMix64; streams 1-8 describe nodes, 16-21 pods), producing the objects the
// reference's own code consumes -- scv.Scv, the pods running on a node, pod label maps.

import (
	"fmt"
	"strconv"

	v1 "k8s.io/api/core/v1"
	metav1 "k8s.io/apimachinery/pkg/apis/meta/v1"

	scv "github.com/NJUPT-ISL/SCV/api/v1"
)

const seed = 0x59304441

// total MB, clock, cores, bandwidth, power -- homogeneous within a node
var skus = [8][5]uint64{{11178, 5705, 3584, 352, 250}, {11019, 7000, 4352, 616, 260}, {16160, 877, 5120, 900, 300}, {32510, 877, 5120, 900, 300},
	{40536, 1215, 6912, 1555, 400}, {81920, 1593, 6912, 2039, 400}, {81559, 2619, 16896, 3350, 700}, {183359, 3996, 18944, 8000, 1000}}

func sm64(x uint64) uint64 {
	z := x + 0x9E3779B97F4A7C15
	z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9
	z = (z ^ (z >> 27)) * 0x94D049BB133111EB
	return z ^ (z >> 31)
}
func rnd(s, stream, i, j uint64) uint64 { return sm64(sm64(sm64(s+stream)+i) + j) }

// synthScv: node idx as an Scv object, the pods running on it (only their scv/memory label matters,
// algorithm.go:75-80), and whether the node has an SCV at all (1 in 256 does not).
func synthScv(idx uint64) (*scv.Scv, []*v1.Pod, bool) {
	sku := skus[rnd(seed, 1, idx, 0)%8]
	odd := rnd(seed, 8, idx, 0)
	ncards := uint64(8)
	if odd%512 == 0 {
		ncards = (odd >> 16) % 9
	}
	obj := &scv.Scv{}
	obj.Name = fmt.Sprintf("node-%d", idx)
	var fsum, tsum uint64
	for g := uint64(0); g < ncards; g++ {
		free := (sku[0] * (rnd(seed, 2, idx, g) >> 32)) >> 32
		health := "Healthy"
		if rnd(seed, 3, idx, g)%64 == 0 {
			health = "Unhealthy"
		}
		obj.Status.CardList = append(obj.Status.CardList, scv.Card{Health: health, FreeMemory: free, TotalMemory: sku[0],
			Clock: uint(sku[1]), Core: uint(sku[2]), Bandwidth: uint(sku[3]), Power: uint(sku[4])})
		fsum, tsum = fsum+free, tsum+sku[0]
	}
	if tsum == 0 {
		tsum = 1 // TotalMemorySum == 0 panics in the reference (algorithm.go:71)
	}
	obj.Status.FreeMemorySum, obj.Status.TotalMemorySum = fsum, tsum
	cn := rnd(seed, 4, idx, 0)
	obj.Status.CardNumber = 8
	if cn%128 == 0 {
		obj.Status.CardNumber = uint((cn >> 8) % 17)
	}
	var running []*v1.Pod
	for k := uint64(0); k < rnd(seed, 6, idx, 0)%9; k++ {
		mem := (rnd(seed, 7, idx, k)%128 + 1) * 256
		running = append(running, &v1.Pod{ObjectMeta: metav1.ObjectMeta{Name: fmt.Sprintf("run-%d-%d", idx, k), Namespace: "default",
			Labels: map[string]string{"scv/memory": strconv.FormatUint(mem, 10)}}})
	}
	return obj, running, rnd(seed, 5, idx, 0)%256 != 0
}

// synthPodLabels: pending pod p of the stream s as its scv/* label map.
func synthPodLabels(p, s uint64) map[string]string {
	l := map[string]string{}
	if rnd(s, 16, p, 0)%100 >= 50 {
		v := rnd(s, 17, p, 0)
		num := []uint64{1, 2, 4, 8}[(v>>8)%4]
		if (v>>20)%100 == 0 {
			num = 9 * (v & 1)
		}
		l["scv/number"] = strconv.FormatUint(num, 10)
	}
	if rnd(s, 18, p, 0)%100 >= 20 {
		l["scv/memory"] = strconv.FormatUint(rnd(s, 19, p, 0)%90000+1, 10)
	}
	if rnd(s, 20, p, 0)%100 >= 85 {
		v := rnd(s, 21, p, 0)
		clk := uint64(1234) // a clock no card has
		if (v>>20)%100 != 0 {
			clk = skus[(v>>8)%8][1]
		}
		l["scv/clock"] = strconv.FormatUint(clk, 10)
	}
	return l
}
