#!/bin/bash
# tools/ncu_heaviest.sh <out-name> <kernel-regex> <cmd...>: (1) plain run, (2) launch list with device times, (3) --set full capture of
# the longest launch matching the regex (index taken from the launch list; the program is deterministic).  Outputs under gpurun_out/.
set -e
name=$1; rx=$2; shift 2
"$@" > gpurun_out/${name}_plain.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${name}_launches.csv "$@" > /dev/null 2>&1
idx=$(python - "$rx" gpurun_out/${name}_launches.csv <<'PY'
import csv, re, sys
rx = re.compile(sys.argv[1]); rows = list(csv.reader(l for l in open(sys.argv[2]) if l.startswith('"')))
h = rows[0]; k = h.index("Kernel Name"); v = h.index("Metric Value")
best = (-1.0, 0); n = 0
for r in rows[1:]:
    if rx.search(r[k]):
        t = float(r[v].replace(",", ""))
        if t > best[0]: best = (t, n)
        n += 1
print(best[1])
PY
)
echo "heaviest launch matching $rx: index $idx" | tee gpurun_out/${name}_pick.txt
ncu --set full --clock-control none --import-source on -k regex:$rx -s $idx -c 1 -o gpurun_out/$name -f "$@" > gpurun_out/${name}_ncu.log 2>&1
ls -la gpurun_out/$name.ncu-rep
