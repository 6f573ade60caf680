// Queue ordering by the scv/priority label (higher first).  Host-side, unchanged in behaviour.
package sort

import (
	"strconv"

	"k8s.io/kubernetes/pkg/scheduler/framework"
)

func Less(a, b *framework.QueuedPodInfo) bool { return GetPodPriority(a) > GetPodPriority(b) }

func GetPodPriority(p *framework.QueuedPodInfo) int {
	if v, ok := p.Pod.Labels["scv/priority"]; ok {
		pri, _ := strconv.Atoi(v)
		return pri
	}
	return 0
}
