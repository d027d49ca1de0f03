python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu13.log 2>&1; echo "gpu rc=$?"; tail -2 gpurun_out/t_gpu13.log
python tools/ab_trace.py c4 "shard8_overlap:shard=8,overlap=1;shard8_nooverlap:shard=8,overlap=0;full_overlap:overlap=1" > gpurun_out/ab13.jsonl 2>&1
grep -o '"workload": "[a-z0-9_]*"\|"options": "[a-z0-9_]*"\|"ms": [0-9.]*\|"trace_ms": [0-9.]*' gpurun_out/ab13.jsonl | paste -sd' ' | sed 's/"workload"/\n"workload"/g'; tail -3 gpurun_out/ab13.jsonl | cut -c1-300
