mkdir -p gpurun_out
timeout 300 python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu21.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/t_gpu21.log
