python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu12.log 2>&1; echo "gpu rc=$?"; tail -2 gpurun_out/t_gpu12.log
python tools/ab_trace.py c2,bunny_box,c3@32,c1 "default:" > gpurun_out/ab11.jsonl 2>&1
grep -o '"workload": "[a-z0-9_]*"\|"ms": [0-9.]*\|"trace_ms": [0-9.]*' gpurun_out/ab11.jsonl | paste -sd' ' | sed 's/"workload"/\n"workload"/g'
python bench.py --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/b9_c4.json 2> gpurun_out/b9_c4.err; python -c "
import json; j=json.loads(open('gpurun_out/b9_c4.json').read().strip().splitlines()[-1]); print('c4', j['value'], j['ms_per_step'], j['roofline']['frac'], j['e2e']['value'])"
