python -m pytest tests/test_multi_gpu.py -x -q -m gpu > gpurun_out/t_multi.log 2>&1; echo "multi rc=$?"; tail -15 gpurun_out/t_multi.log
python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu5.log 2>&1; echo "gpu rc=$?"; tail -3 gpurun_out/t_gpu5.log
