"""Every COLLADA scene the reference ships, loaded by the reference's own loader (oracle/_ref/ref_probe scene: collada parser, half-edge mesh,
Application::init_*) and by host/dae_loader.cpp (host/pathtracer --dump-scene): the two flattened scenes must be the same bytes.
Runs only where /root/reference is present (the build container).  usage: compare_loaders.py [dir with .dae files]"""
import glob, os, subprocess, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import importlib
import numpy as np
sc = importlib.import_module("relativistic-ray-tracer_b200.scene")
src = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/pathtracer/dae"
tmp = tempfile.mkdtemp(prefix="rrt_cmp_")
same = diff = 0
for dae in sorted(glob.glob(os.path.join(src, "**", "*.dae"), recursive=True)):
    name = os.path.relpath(dae, src)
    a, b = os.path.join(tmp, "ref.rrtscene"), os.path.join(tmp, "host.rrtscene")
    for f in (a, b):
        if os.path.exists(f):
            os.remove(f)
    r1 = subprocess.run([os.path.join(ROOT, "oracle", "_ref", "ref_probe"), "scene", dae, "480", "360", a], capture_output=True, text=True, timeout=600)
    r2 = subprocess.run([os.path.join(ROOT, "host", "pathtracer"), "-r", "480", "360", "--dump-scene", b, dae], capture_output=True, text=True, timeout=600)
    ok1, ok2 = os.path.exists(a), os.path.exists(b)
    if not ok1 or not ok2:
        print("%-40s reference loader %s, host loader %s   %s" % (name, "ok" if ok1 else "FAILED rc=%d" % r1.returncode, "ok" if ok2 else "refused",
                                                                (r2.stderr.strip().splitlines() or [""])[0][:120] if not ok2 else ""))
        same += (not ok1 and not ok2); diff += (ok1 != ok2)
        continue
    ra, rb = open(a, "rb").read(), open(b, "rb").read()
    if ra == rb:
        s = sc.SceneData.load(a)
        print("%-40s identical: %d triangles, %d spheres, %d materials, %d lights" % (name, s.tri_pos.shape[0], s.sph.shape[0], len(s.mats), len(s.lights)))
        same += 1
    else:
        sa, sb = sc.SceneData.load(a), sc.SceneData.load(b)
        what = [k for k in ("prim_kind", "prim_index", "tri_pos", "tri_nrm", "tri_mat", "sph", "sph_mat", "mats", "lights")
                if getattr(sa, k).shape != getattr(sb, k).shape or getattr(sa, k).tobytes() != getattr(sb, k).tobytes()]
        cam = [k for k in sa.camera if sa.camera[k] != sb.camera[k]]
        print("%-40s DIFFERENT: %s %s" % (name, what, cam))
        diff += 1
print("%d scenes agree, %d differ" % (same, diff))
