mkdir -p gpurun_out; rm -f gpurun_out/ab24.jsonl
timeout 200 python tools/ab_trace.py c4 "base:;in8:inflight=8388608,workgb=64;in16:inflight=16777216,workgb=64;in2:inflight=2097152" >> gpurun_out/ab24.jsonl 2>&1 || echo "FAILED rc=$?" >> gpurun_out/ab24.jsonl
grep -o 'FAILED.*\|Error.*\|"options": "[a-z0-9_]*"\|"ms": [0-9.]*\|"trace_ms": [0-9.]*\|"sha": "[0-9a-f]*"' gpurun_out/ab24.jsonl | paste -sd' ' | sed 's/"options"/\n"options"/g'
