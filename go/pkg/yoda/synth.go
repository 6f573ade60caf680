package yoda

import "github.com/NJUPT-ISL/Yoda-Scheduler/pkg/yoda/b200"

// synthNode / synthPod: Go twins of yoda_scheduler_b200/synth.py (streams 1-8 nodes, 16-21 pods).
func synthNode(idx uint64) (*b200.NodeRecord, uint64, bool) {
	sku := skus[rnd(seed, 1, idx, 0)%8]
	odd := rnd(seed, 8, idx, 0)
	ncards := uint64(8)
	if odd%512 == 0 {
		ncards = (odd >> 16) % 9
	}
	free, total, healthy := make([]uint64, ncards), make([]uint64, ncards), uint32(0)
	for g := uint64(0); g < ncards; g++ {
		total[g] = sku[0]
		free[g] = (sku[0] * (rnd(seed, 2, idx, g) >> 32)) >> 32
		if rnd(seed, 3, idx, g)%64 != 0 {
			healthy |= 1 << g
		}
	}
	cn := rnd(seed, 4, idx, 0)
	cardNumber := uint64(8)
	if cn%128 == 0 {
		cardNumber = (cn >> 8) % 17
	}
	var alloc uint64
	for k := uint64(0); k < rnd(seed, 6, idx, 0)%9; k++ {
		alloc += (rnd(seed, 7, idx, k)%128 + 1) * 256
	}
	return b200.MakeRecord(cardNumber, healthy, free, total, sku[1], sku[2], sku[3], sku[4]), alloc, rnd(seed, 5, idx, 0)%256 != 0
}

func synthPod(p, s uint64) b200.PodReq {
	var hasNum, hasMem, hasClk bool
	var num, mem, clk uint64
	if rnd(s, 16, p, 0)%100 >= 50 {
		hasNum = true
		v := rnd(s, 17, p, 0)
		if (v>>20)%100 == 0 {
			num = 9 * (v & 1)
		} else {
			num = []uint64{1, 2, 4, 8}[(v>>8)%4]
		}
	}
	if rnd(s, 18, p, 0)%100 >= 20 {
		hasMem, mem = true, rnd(s, 19, p, 0)%90000+1
	}
	if rnd(s, 20, p, 0)%100 >= 85 {
		hasClk = true
		v := rnd(s, 21, p, 0)
		clk = 1234
		if (v>>20)%100 != 0 {
			clk = skus[(v>>8)%8][1]
		}
	}
	return b200.MakeReq(hasNum, hasMem, hasClk, num, mem, clk)
}
