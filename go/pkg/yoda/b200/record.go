package b200

/*
#include "yoda_b200.h"
*/
import "C"

import scv "github.com/NJUPT-ISL/SCV/api/v1"

// RecordFromScv flattens the fields of an Scv object that the hot path reads.
func RecordFromScv(s *scv.Scv) *NodeRecord {
	var r NodeRecord
	r.card_number = C.uint64_t(s.Status.CardNumber)
	r.free_memory_sum = C.uint64_t(s.Status.FreeMemorySum)
	r.total_memory_sum = C.uint64_t(s.Status.TotalMemorySum)
	n := len(s.Status.CardList)
	if n > MaxCards {
		n = MaxCards + 1 // rejected by the library with YODA_E_RANGE
	}
	r.n_cards = C.uint32_t(n)
	for g, c := range s.Status.CardList {
		if g >= MaxCards {
			break
		}
		if c.Health == "Healthy" {
			r.healthy_mask |= 1 << uint(g)
		}
		r.free_memory[g] = C.uint64_t(c.FreeMemory)
		r.total_memory[g] = C.uint64_t(c.TotalMemory)
		r.clock[g] = C.uint64_t(c.Clock)
		r.core[g] = C.uint64_t(c.Core)
		r.bandwidth[g] = C.uint64_t(c.Bandwidth)
		r.power[g] = C.uint64_t(c.Power)
	}
	return &r
}
