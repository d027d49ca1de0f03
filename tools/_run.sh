mkdir -p gpurun_out
timeout 400 python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu19.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/t_gpu19.log
