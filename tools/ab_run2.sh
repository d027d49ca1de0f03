#!/bin/bash
# tools/ab_run2.sh "bench args" lib...   -> one line per library: ms per frame, Mpaths/s, trace ms
a=$1; shift
for lib in "$@"; do
  RRT_B200_LIB=$lib python bench.py $a --steps 4 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$lib', '$a', round(d['ms_per_step'],3), round(d['value'],2), round(d['roofline']['trace_ms_per_step'],3))"
done
