python -m pytest tests/test_multi_gpu.py -x -q -m gpu 2>&1 | tail -3
python -c "
import importlib, sys
sys.path.insert(0,'tests')
api=importlib.import_module('relativistic-ray-tracer_b200.api')
m=api.Context([0,1]); print('multi ctx ok', m.gpu_count())
import torch; print('torch imported after rrt_create_multi:', torch.cuda.device_count(), 'GPUs')
" 2>&1 | tail -3
host/pathtracer -s 4 -l 2 -m 3 -G 0,1 -f gpurun_out/cli_multi.png tests/golden/cbspheres_lambertian_480x360.rrtscene 2>&1 | tail -2
