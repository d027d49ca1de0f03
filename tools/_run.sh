python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu16.log 2>&1; echo "gpu rc=$?"; tail -2 gpurun_out/t_gpu16.log
python tools/ab_trace.py bunny_box,bunny_bh,c5,c3@32 "default:" > gpurun_out/ab16.jsonl 2>&1
grep -o '"workload": "[a-z0-9_]*"\|"ms": [0-9.]*\|"trace_ms": [0-9.]*\|"tri_records": [0-9]*' gpurun_out/ab16.jsonl | paste -sd' ' | sed 's/"workload"/\n"workload"/g'
