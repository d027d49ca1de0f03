python -m pytest tests -x -q -m gpu -s > gpurun_out/t_gpu10.log 2>&1; echo "gpu rc=$?"; grep "full-size\|grazing\|differing\|independent\|passed\|failed\|Error" gpurun_out/t_gpu10.log | tail -14
python tools/ab_trace.py c5 "top:top=1;notop:top=0" > gpurun_out/ab8.jsonl 2>&1
grep -o '"workload": "[a-z0-9_]*"\|"options": "[a-z_]*"\|"ms": [0-9.]*\|"trace_ms": [0-9.]*\|"nodes": [0-9]*' gpurun_out/ab8.jsonl | paste -sd' ' | sed 's/"workload"/\n"workload"/g'
python bench.py --steps 3 --warmup 2 > gpurun_out/b5_c4.json 2> gpurun_out/b5_c4.err; python -c "
import json; j=json.loads(open('gpurun_out/b5_c4.json').read().strip().splitlines()[-1]); print('c4', j['value'], j['ms_per_step'], j['roofline']['frac'], j['e2e']['value'], j['cpu_baseline']['value'])"
