mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_multi_gpu.py -x -q -m gpu > gpurun_out/t_multi3.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/t_multi3.log
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/final3_c4_2gpu.json 2> gpurun_out/final3_c4_2gpu.err; echo "bench rc=$?"; cut -c1-260 gpurun_out/final3_c4_2gpu.json
