// Binary entry: the kube-scheduler command with the yoda plugin registered, plus a `bench`
// sub-command that drives the plugin through a fake framework handle (no cluster needed).
package main

import (
	"fmt"
	"math/rand"
	"os"
	"time"

	"github.com/NJUPT-ISL/Yoda-Scheduler/pkg/register"
	"github.com/NJUPT-ISL/Yoda-Scheduler/pkg/yoda"
	"k8s.io/component-base/logs"
)

func main() {
	if len(os.Args) > 1 && os.Args[1] == "bench" {
		os.Exit(yoda.RunBench(os.Args[2:]))
	}
	rand.Seed(time.Now().UTC().UnixNano())
	command := register.Register()
	logs.InitLogs()
	defer logs.FlushLogs()
	if err := command.Execute(); err != nil {
		_, _ = fmt.Fprintf(os.Stderr, "%v\n", err)
		os.Exit(1)
	}
}
