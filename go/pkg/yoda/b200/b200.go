// Package b200 is the thin cgo shim over include/yoda_b200.h.  No Go pointer is retained by C:
// every buffer is passed for the duration of one call only.
package b200

// Search paths come from the environment, because this package lives inside a copy of the reference tree
// (see go/apply_overlay.sh): CGO_CFLAGS=-I<repo>/include CGO_LDFLAGS="-L<repo>/yoda_scheduler_b200 -Wl,-rpath,...".

/*
#cgo LDFLAGS: -lyoda_b200
#include <stdlib.h>
#include "yoda_b200.h"
*/
import "C"

import (
	"fmt"
	"runtime"
	"unsafe"
)

const (
	MaxCards    = C.YODA_MAX_CARDS
	HasNumber   = C.YODA_POD_HAS_NUMBER
	HasMemory   = C.YODA_POD_HAS_MEMORY
	HasClock    = C.YODA_POD_HAS_CLOCK
	RowFeasible = C.YODA_ROW_FEASIBLE_BITS
	RowRaw      = C.YODA_ROW_RAW
	RowNorm     = C.YODA_ROW_NORM
)

type (
	NodeRecord = C.yoda_node_record
	PodReq     = C.yoda_pod_req
	PodResult  = C.yoda_pod_result
)

type Ctx struct{ h *C.yoda_ctx }

// call runs one C entry point and, if it failed, fetches its message.  yoda_last_error() is thread-local on
// the C side and a goroutine may be moved to another OS thread between two cgo calls, so the pair is made
// on a locked thread: the message read is the one the failing call wrote.
func call(f func() C.int) error {
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	if rc := f(); rc != 0 {
		return fmt.Errorf("yoda_b200 error %d: %s", int(rc), C.GoString(C.yoda_last_error()))
	}
	return nil
}

func New(device int, maxNodes uint32) (*Ctx, error) {
	cfg := C.yoda_config{device: C.int32_t(device), max_nodes: C.uint32_t(maxNodes), shard_count: 1}
	var h *C.yoda_ctx
	if err := call(func() C.int { return C.yoda_create(&cfg, &h) }); err != nil {
		return nil, err
	}
	return &Ctx{h}, nil
}

func (c *Ctx) Close() { C.yoda_destroy(c.h); c.h = nil }

func (c *Ctx) Upsert(idx uint32, rec *NodeRecord) error {
	return call(func() C.int { return C.yoda_node_upsert(c.h, C.uint32_t(idx), rec) })
}

func (c *Ctx) SetAlloc(idx uint32, sum uint64) error {
	return call(func() C.int { return C.yoda_node_set_alloc(c.h, C.uint32_t(idx), C.uint64_t(sum)) })
}

func (c *Ctx) Remove(idx uint32) error {
	return call(func() C.int { return C.yoda_node_remove(c.h, C.uint32_t(idx)) })
}

func (c *Ctx) Commit() (uint64, error) {
	var e C.uint64_t
	if err := call(func() C.int { return C.yoda_commit(c.h, &e) }); err != nil {
		return 0, err
	}
	return uint64(e), nil
}

func (c *Ctx) NodeCount() uint32 {
	var n C.uint32_t
	C.yoda_node_count(c.h, &n)
	return uint32(n)
}

// Eval: Filter + CollectMaxValues + Score + NormalizeScore + node choice for all pods in one pass.
func (c *Ctx) Eval(pods []PodReq, full bool) ([]PodResult, error) {
	out := make([]PodResult, len(pods))
	if len(pods) == 0 {
		return out, nil
	}
	flags := C.uint32_t(C.YODA_EVAL_DECIDE)
	if full {
		// rows are only ever read through yoda_fetch_row (int64 either way): let the device keep the scores as uint32
		flags = C.uint32_t(C.YODA_EVAL_FULL | C.YODA_EVAL_RAW32)
	}
	err := call(func() C.int {
		return C.yoda_eval_batch(c.h, (*C.yoda_pod_req)(unsafe.Pointer(&pods[0])), C.uint32_t(len(pods)), flags,
			(*C.yoda_pod_result)(unsafe.Pointer(&out[0])))
	})
	if err != nil {
		return nil, err
	}
	return out, nil
}

func (c *Ctx) Row(pod uint32, kind int, dst unsafe.Pointer, bytes int) error {
	return call(func() C.int { return C.yoda_fetch_row(c.h, C.uint32_t(pod), C.int(kind), dst, C.size_t(bytes)) })
}

// NormalizeScores: scheduler.go:132-157 on an arbitrary list, in place, on the device.
func (c *Ctx) NormalizeScores(scores []int64) error {
	return call(func() C.int {
		return C.yoda_normalize_scores(c.h, (*C.int64_t)(unsafe.Pointer(&scores[0])), C.uint32_t(len(scores)))
	})
}

// QueueOrder: batch order under sort.Less (scv/priority, higher first, stable).
func (c *Ctx) QueueOrder(priority []int64) ([]uint32, error) {
	order := make([]uint32, len(priority))
	if len(priority) == 0 {
		return order, nil
	}
	err := call(func() C.int {
		return C.yoda_queue_order(c.h, (*C.int64_t)(unsafe.Pointer(&priority[0])), C.uint32_t(len(priority)), (*C.uint32_t)(unsafe.Pointer(&order[0])))
	})
	if err != nil {
		return nil, err
	}
	return order, nil
}

// EvalSequential: pods in `order`, each choice added to the chosen node's allocateMemorySum before the next pod.
func (c *Ctx) EvalSequential(pods []PodReq, order []uint32) ([]PodResult, error) {
	out := make([]PodResult, len(pods))
	if len(pods) == 0 {
		return out, nil
	}
	var op *C.uint32_t
	if order != nil {
		op = (*C.uint32_t)(unsafe.Pointer(&order[0]))
	}
	err := call(func() C.int {
		return C.yoda_eval_sequential(c.h, (*C.yoda_pod_req)(unsafe.Pointer(&pods[0])), C.uint32_t(len(pods)), op, (*C.yoda_pod_result)(unsafe.Pointer(&out[0])))
	})
	if err != nil {
		return nil, err
	}
	return out, nil
}

// MakeRecord / MakeReq / Chosen: helpers for callers outside this package (cgo field names are
// not exported across packages).
func MakeRecord(cardNumber uint64, healthy uint32, free, total []uint64, clock, core, bandwidth, power uint64) *NodeRecord {
	var r NodeRecord
	r.card_number = C.uint64_t(cardNumber)
	r.n_cards = C.uint32_t(len(free))
	r.healthy_mask = C.uint32_t(healthy)
	var fs, ts uint64
	for g := range free {
		r.free_memory[g], r.total_memory[g] = C.uint64_t(free[g]), C.uint64_t(total[g])
		r.clock[g], r.core[g], r.bandwidth[g], r.power[g] = C.uint64_t(clock), C.uint64_t(core), C.uint64_t(bandwidth), C.uint64_t(power)
		fs, ts = fs+free[g], ts+total[g]
	}
	if ts == 0 {
		ts = 1
	}
	r.free_memory_sum, r.total_memory_sum = C.uint64_t(fs), C.uint64_t(ts)
	return &r
}

func MakeReq(hasNum, hasMem, hasClk bool, num, mem, clk uint64) PodReq {
	var r PodReq
	if hasNum {
		r.flags |= HasNumber
		r.number = C.uint64_t(num)
	}
	if hasMem {
		r.flags |= HasMemory
		r.memory = C.uint64_t(mem)
	}
	if hasClk {
		r.flags |= HasClock
		r.clock = C.uint64_t(clk)
	}
	return r
}

func Chosen(r *PodResult) int32    { return int32(r.chosen) }
func Feasible(r *PodResult) uint32 { return uint32(r.feasible) }
func Maxima(r *PodResult) [6]uint64 {
	return [6]uint64{uint64(r.max_bandwidth), uint64(r.max_clock), uint64(r.max_core), uint64(r.max_free_memory), uint64(r.max_power), uint64(r.max_total_memory)}
}
