#!/bin/bash
# wall-clock of host/pathtracer on the C5 .dae at two frame sizes, of the load alone, and on a tiny scene at 4K / 480x360
python - <<'PY'
import importlib,os,sys
sys.path.insert(0,"."); sys.path.insert(0,"tests")
synth = importlib.import_module("relativistic-ray-tracer_b200.synth")
os.makedirs("/tmp/x",exist_ok=True)
synth.write_c5_dae("/tmp/x/c5.dae", open("oracle/_ref/dae/CBspheres_lambertian.dae").read())
PY
t() { local s=$(date +%s%N); "$@" > /dev/null 2>&1; local e=$(date +%s%N); echo "$(( (e - s) / 1000000 )) ms  $*"; }
for r in "480 360" "3840 2160"; do for i in 1 2; do t host/pathtracer -s 4 -l 4 -m 5 -r $r -f /tmp/x/o.png /tmp/x/c5.dae; done; done
t host/pathtracer --dump-scene /dev/null /tmp/x/c5.dae
t host/pathtracer -s 4 -l 4 -m 5 -r 3840 2160 -f /tmp/x/o2.png tests/golden/cbspheres_lambertian_480x360.rrtscene
t host/pathtracer -s 4 -l 4 -m 5 -r 480 360 -f /tmp/x/o3.png tests/golden/cbspheres_lambertian_480x360.rrtscene
