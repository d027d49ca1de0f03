python -m pytest tests -x -q -m gpu -s > gpurun_out/t_gpu6.log 2>&1; echo "gpu rc=$?"; grep "full-size\|passed\|failed\|Error\|assert" gpurun_out/t_gpu6.log | tail -12
python tools/ab_trace.py bunny_bh,c5,c2,bunny_box "escape:escape=1;noescape:escape=0" > gpurun_out/ab4.jsonl 2>&1; cut -c1-150 gpurun_out/ab4.jsonl
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/b4_c4.json 2> gpurun_out/b4_c4.err; python -c "
import json; j=json.loads(open('gpurun_out/b4_c4.json').read().strip().splitlines()[-1]); print('c4', j['value'], j['ms_per_step'], j['roofline']['frac'])"
