module github.com/NJUPT-ISL/Yoda-Scheduler

go 1.15

require (
	github.com/NJUPT-ISL/SCV v0.0.0-20200901022803-46b36eeed646
	github.com/spf13/cobra v1.1.1
	k8s.io/api v0.20.0
	k8s.io/apimachinery v0.20.0
	k8s.io/client-go v0.20.0
	k8s.io/component-base v0.20.0
	k8s.io/klog v1.0.0
	k8s.io/kubernetes v1.20.0
	sigs.k8s.io/controller-runtime v0.7.0
)
