for lib in relativistic-ray-tracer_b200/csrc/librrt_b200.so build_ab/lib_pool16.so build_ab/lib_pool64.so; do
echo "== $lib" >> gpurun_out/ab18.jsonl
RRT_B200_LIB=$lib python tools/ab_trace.py c4 "shard8:shard=8" >> gpurun_out/ab18.jsonl 2>&1
done
grep -o '== .*\|"ms": [0-9.]*\|"trace_ms": [0-9.]*' gpurun_out/ab18.jsonl | paste -sd' ' | sed 's/== /\n== /g'
