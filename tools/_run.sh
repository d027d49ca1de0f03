python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "oracle_records or bvh_intersect or full_size or device_bvh or far_origin" > gpurun_out/t_gpu14.log 2>&1; echo "gpu rc=$?"; tail -2 gpurun_out/t_gpu14.log
python tools/ab_trace.py bunny_box,bunny_bh,c5,c3@32 "default:" > gpurun_out/ab14.jsonl 2>&1
grep -o '"workload": "[a-z0-9_]*"\|"ms": [0-9.]*\|"trace_ms": [0-9.]*\|"nodes": [0-9]*' gpurun_out/ab14.jsonl | paste -sd' ' | sed 's/"workload"/\n"workload"/g'
