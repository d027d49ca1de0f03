python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu17.log 2>&1; echo "gpu rc=$?"; tail -2 gpurun_out/t_gpu17.log
bash tools/ncu_heaviest.sh r02_c4_shadow_final 'k_trace_shadow' python tools/one_frame.py c4 fp64 1; tail -1 gpurun_out/r02_c4_shadow_final_plain.log | cut -c200-
