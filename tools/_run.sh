python -m pytest tests/test_host_gpu.py tests/test_multi_gpu.py -x -q -m gpu 2>&1 | tail -2
python - <<'PY'
import importlib,os,sys
sys.path.insert(0,"."); sys.path.insert(0,"tests")
synth = importlib.import_module("relativistic-ray-tracer_b200.synth")
os.makedirs("/tmp/x",exist_ok=True)
synth.write_c5_dae("/tmp/x/c5.dae", open("oracle/_ref/dae/CBspheres_lambertian.dae").read())
PY
S=tests/golden/cbspheres_lambertian_480x360.rrtscene
{
for i in 1 2; do s=$(date +%s%N); RRT_HOST_TIMING=1 host/pathtracer -s 16 -l 4 -m 5 -r 480 360 -f gpurun_out/cli_c2.png $S 2>&1 | grep "timing"; e=$(date +%s%N); echo "$(( (e - s) / 1000000 )) ms wall: C2 command line"; done
for i in 1 2; do s=$(date +%s%N); RRT_HOST_TIMING=1 host/pathtracer -s 4 -l 4 -m 5 -r 3840 2160 -f gpurun_out/cli_c5.png /tmp/x/c5.dae 2>&1 | grep "timing"; e=$(date +%s%N); echo "$(( (e - s) / 1000000 )) ms wall: C5 .dae (52 MB, 1 M triangles) at 4K / 4 spp"; done
} > gpurun_out/cli_timing2.txt 2>&1
cat gpurun_out/cli_timing2.txt
rm -f gpurun_out/cli_c5.png gpurun_out/cli_c5_rate.png
