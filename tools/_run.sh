python bench.py > gpurun_out/final_c4.json 2> gpurun_out/final_c4.err; echo "c4 rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/final_c4_ref.json 2> gpurun_out/final_c4_ref.err; echo "ref rc=$?"
python bench.py --workload c5 --spp 64 --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/final_c5_64spp.json 2> gpurun_out/final_c5.err; echo "c5 rc=$?"
python bench.py --workload c2 --steps 5 --warmup 3 > gpurun_out/final_c2.json 2> gpurun_out/final_c2.err; echo "c2 rc=$?"
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
