python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu3.log 2>&1; echo "gpu rc=$?"; tail -2 gpurun_out/t_gpu3.log
python bench.py > gpurun_out/b_c4.json 2> gpurun_out/b_c4.err; echo "bench rc=$?"; tail -c 400 gpurun_out/b_c4.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/b_c4_ref.json 2> gpurun_out/b_c4_ref.err; echo "ref rc=$?"
python bench.py --workload c2 --steps 5 > gpurun_out/b_c2.json 2> gpurun_out/b_c2.err; echo "c2 rc=$?"
