// Package collection keeps the CycleState contract of the plugin: key "Max" holds *Data.
// The maxima themselves are computed on the device (see csrc/fast_kernels.cuh, k_max_*).
package collection

import "k8s.io/kubernetes/pkg/scheduler/framework"

type MaxValue struct {
	MaxBandwidth   uint
	MaxClock       uint
	MaxCore        uint
	MaxFreeMemory  uint64
	MaxPower       uint
	MaxTotalMemory uint64
}

type Data struct{ Value MaxValue }

func (d *Data) Clone() framework.StateData { return &Data{Value: d.Value} }
