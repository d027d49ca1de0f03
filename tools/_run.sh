for n in 8 4 2; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/b10_c4_${n}gpu.json 2> gpurun_out/b10_c4_${n}gpu.err; echo "n=$n rc=$?"
done
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/b10_c4_1gpu.json 2> gpurun_out/b10_c4_1gpu.err; echo "n=1 rc=$?"
python -m pytest tests/test_multi_gpu.py -x -q -m gpu 2>&1 | tail -1
python - <<'PY' > gpurun_out/multi_in_library.txt 2>&1
# the same frame through ONE process driving the GPUs inside the library (rrt_create_multi)
import importlib, sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import bench
from util import sc
api = importlib.import_module("relativistic-ray-tracer_b200.api")
s = bench.workload_scene("c4")
for devs in ([0], [0, 1], [0, 1, 2, 3], list(range(8))):
    c = api.Context(devs if len(devs) > 1 else devs[0]); c.upload_scene(s); c.set_frame(1920, 1080); c.set_option(api.OPT_COUNTERS, 0)
    p = sc.make_params(ns_aa=1024, max_ray_depth=5, ns_area_light=4, samples_per_batch=32, max_tolerance=0.05, seed=1)
    c.render(p)
    t0 = time.perf_counter(); hdr, cnt, st = c.render(p); dt = time.perf_counter() - t0
    print("rrt_create_multi on %d GPU(s): %.1f ms wall per rrt_render (host buffers), %.1f ms device, %.1f Mpaths/s" % (len(devs), dt * 1e3, st["seconds"] * 1e3, cnt.sum() / dt / 1e6), flush=True)
    c.close()
PY
cat gpurun_out/multi_in_library.txt
