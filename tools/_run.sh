run8() { # name lib streams
RRT_B200_LIB=$2 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 2 --warmup 1 --no-cpu-baseline --streams $3 > gpurun_out/b3_$1.json 2> gpurun_out/b3_$1.err; echo "$1 rc=$?"
}
D=relativistic-ray-tracer_b200/csrc/librrt_b200.so
run8 p128_s2 $D 2
run8 p128_s4 $D 4
run8 p32_s2 build_ab/lib_pool32.so 2
run8 p32_s4 build_ab/lib_pool32.so 4
run8 p64_s3 build_ab/lib_pool64.so 3
run8 p32_s1 build_ab/lib_pool32.so 1
for lib in $D build_ab/lib_pool32.so build_ab/lib_pool64.so; do
RRT_B200_LIB=$lib python tools/ab_trace.py bunny_box,c2,bunny_bh,c5 "s2:streams=2;s4:streams=4" >> gpurun_out/ab3.jsonl 2>&1
done
