"""The reference's command line on the C5 scene, end to end on a GPU box: export the procedural scene as COLLADA, run host/pathtracer
on it (load + flatten + BVH build on the device + render + PNG), and check the PNG against the Python binding's render of the scene
file the CLI dumped.  Prints wall-clock times."""
import importlib, os, subprocess, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from util import sc
from test_host_gpu import read_png
api = importlib.import_module("relativistic-ray-tracer_b200.api")
synth = importlib.import_module("relativistic-ray-tracer_b200.synth")
tmp = "/tmp/rrt_c5_cli"; os.makedirs(tmp, exist_ok=True)
tpl = open(os.path.join(ROOT, "oracle", "_ref", "dae", "CBspheres_lambertian.dae")).read()
t = time.perf_counter(); synth.write_c5_dae(tmp + "/c5.dae", tpl); print("export .dae      %.2f s (%.0f MB)" % (time.perf_counter() - t, os.path.getsize(tmp + "/c5.dae") / 1e6))
exe = os.path.join(ROOT, "host", "pathtracer")
W, H, spp = 3840, 2160, int(sys.argv[1]) if len(sys.argv) > 1 else 4
for k in range(2):
    t = time.perf_counter()
    r = subprocess.run([exe, "-s", str(spp), "-l", "4", "-m", "5", "-r", str(W), str(H), "-S", "1", "-f", tmp + "/c5.png", tmp + "/c5.dae"], capture_output=True, text=True)
    wall = time.perf_counter() - t
    assert r.returncode == 0, r.stdout + r.stderr
    print("host/pathtracer  %.2f s wall (%s run): %dx%d, %d spp = %.1f M paths" % (wall, "cold" if k == 0 else "warm", W, H, spp, W * H * spp / 1e6))
for line in r.stdout.replace("\r", "\n").split("\n"):
    if "Done" in line or "BVH traced" in line or "100%" in line:
        print("   ", line.strip()[-110:])
subprocess.run([exe, "--dump-scene", tmp + "/c5.rrtscene", "-r", str(W), str(H), tmp + "/c5.dae"], check=True, capture_output=True)
s = sc.SceneData.load(tmp + "/c5.rrtscene")
c = api.Context(0); c.upload_scene(s); c.set_frame(W, H)
hdr, cnt, st = c.render(sc.make_params(ns_aa=spp, max_ray_depth=5, ns_area_light=4, seed=1))
w = np.ascontiguousarray(c.tonemap(hdr).reshape(H, W)[::-1]).view(np.uint8).reshape(H, W, 4).copy(); w[..., 3] = 255
png = read_png(tmp + "/c5.png")
print("PNG equals the binding's render:", bool(np.array_equal(png, w)), " render %.1f ms, %.0f Mpaths/s" % (st["seconds"] * 1e3, st["paths"] / st["seconds"] / 1e6))
