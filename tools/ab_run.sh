#!/bin/bash
# tools/ab_run.sh workload lib...   -> one line per library: ms per frame, Mpaths/s, trace ms
w=$1; shift
for lib in "$@"; do
  RRT_B200_LIB=$lib python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$lib', '$w', round(d['ms_per_step'],3), round(d['value'],2), round(d['roofline']['trace_ms_per_step'],3))"
done
