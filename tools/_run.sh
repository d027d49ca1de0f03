python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu15.log 2>&1; echo "gpu rc=$?"; tail -2 gpurun_out/t_gpu15.log
python bench.py --steps 3 --warmup 2 > gpurun_out/final2_c4.json 2> gpurun_out/final2_c4.err; python -c "
import json; j=json.loads(open('gpurun_out/final2_c4.json').read().strip().splitlines()[-1]); print('c4', j['value'], j['ms_per_step'], j['roofline']['frac'], j['e2e']['value'], j['cpu_baseline']['value'])"
python tools/ab_trace.py c4 "shard8:shard=8" 2>&1 | head -1 | cut -c1-120
