#!/bin/sh
# apply_overlay.sh <reference checkout> <destination> [repo root]
# Builds the drop-in tree: a copy of the reference (github.com/NJUPT-ISL/Yoda-Scheduler) with this directory
# laid over it.  The reference's pkg/register, pkg/yoda/filter, pkg/yoda/collection, pkg/yoda/score and
# pkg/yoda/sort stay as they are; pkg/yoda/scheduler.go, cmd/scheduler/main.go and go.mod are replaced;
# pkg/yoda/b200 (cgo shim), pkg/yoda/bench.go and pkg/yoda/synth.go are added.
set -eu
REF=$1
DST=$2
REPO=${3:-$(cd "$(dirname "$0")/.." && pwd)}
mkdir -p "$DST"
cp -R "$REF"/. "$DST"/
cp -R "$(dirname "$0")"/cmd "$(dirname "$0")"/pkg "$(dirname "$0")"/go.mod "$DST"/
rm -f "$DST"/go.sum   # regenerate: go mod tidy (needs the module proxy or a filled module cache)
echo "overlay written to $DST"
echo "build:  cd $DST && CGO_CFLAGS=-I$REPO/include CGO_LDFLAGS='-L$REPO/yoda_scheduler_b200 -lyoda_b200 -Wl,-rpath,$REPO/yoda_scheduler_b200' go build ./cmd/scheduler"
echo "golden: cp $REPO/go/refgolden/golden_test.go $REF/pkg/yoda/zz_golden_test.go && (cd $REF && go test ./pkg/yoda -run TestGolden -args -golden $REPO/tests/golden)"
