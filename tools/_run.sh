mkdir -p gpurun_out; rm -f gpurun_out/ab19.jsonl
for lib in build_ab/lib_noew.so relativistic-ray-tracer_b200/csrc/librrt_b200.so; do
echo "== $lib" >> gpurun_out/ab19.jsonl
RRT_B200_LIB=$lib timeout 150 python tools/ab_trace.py bunny_box,c2,c4 >> gpurun_out/ab19.jsonl 2>&1 || echo "FAILED rc=$?" >> gpurun_out/ab19.jsonl
done
grep -o '== .*\|FAILED.*\|"workload": "[a-z0-9_]*"\|"ms": [0-9.]*\|"trace_ms": [0-9.]*\|"sha": "[0-9a-f]*"' gpurun_out/ab19.jsonl | paste -sd' ' | sed 's/== /\n== /g'
