#!/bin/bash
# run bench.py phases for every hand-built variant library yoda_scheduler_b200/libyoda_var_*.so (YODA_B200_SO override)
for so in yoda_scheduler_b200/libyoda_b200.so yoda_scheduler_b200/libyoda_var_*.so; do
  for shape in "--config c3" "--config c3 --raw32" "--config c3 --nodes 125000 --pods 10000" "--config c5"; do
    YODA_B200_SO=$PWD/$so python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-latency --no-c4 $shape 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('$so', '$shape', 'ms', round(d['ms_per_step'],3), {k: round(v,3) for k,v in d['phases_ms'].items()})"
  done
done
