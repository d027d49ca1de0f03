mkdir -p gpurun_out
timeout 200 python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu22.log 2>&1; echo "pytest rc=$?"; tail -n 2 gpurun_out/t_gpu22.log
