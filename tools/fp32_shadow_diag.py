import importlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from util import golden, load_scene, sc, hits_from_probe
api = importlib.import_module("relativistic-ray-tracer_b200.api")
np.set_printoptions(linewidth=220, precision=9)
scene = "cbspheres_lambertian_480x360"
S = load_scene(scene)
L = S.lights[0]
print("light", L)
c0 = api.Context(0); c0.upload_scene(S)
c1 = api.Context(0); c1.set_option(api.OPT_STEPPER, 1); c1.upload_scene(S)
rng = np.random.default_rng(5)
xy = rng.random((200000, 2))
ro, rd = c0.generate_rays(xy)
h0, _ = (c1 if len(sys.argv) > 1 else c0).trace_rays(ro, rd)
m = h0["status"] == 1
p = h0["hit_p"][m]
xi = rng.random((len(p), 2)) - 0.5
q = np.asarray(L["position"]) + xi[:, :1] * np.asarray(L["dim_x"]) + xi[:, 1:] * np.asarray(L["dim_y"])
wi = q - p; wi /= np.linalg.norm(wi, axis=1, keepdims=True)
o = p + 1e-11 * wi
a0, _ = c0.trace_rays(o, wi, any_hit=True); a1, _ = c1.trace_rays(o, wi, any_hit=True)
f0, _ = c0.trace_rays(o, wi); f1, _ = c1.trace_rays(o, wi)
print("shadow rays", len(o), "any status fp64", np.bincount(a0["status"], minlength=3), "fp32", np.bincount(a1["status"], minlength=3),
      "mismatch", float((a0["status"] != a1["status"]).mean()), "steps", a0["steps"].sum(), a1["steps"].sum())
print("closest status fp64", np.bincount(f0["status"], minlength=3), "fp32", np.bincount(f1["status"], minlength=3), "mismatch",
      float(((f0["status"] != f1["status"]) | (f0["prim"] != f1["prim"])).mean()))
bad = np.where((f0["status"] != f1["status"]) | (f0["prim"] != f1["prim"]))[0]
from collections import Counter
print("mismatch kinds (fp64 status, prim -> fp32 status, prim):", Counter((int(f0["status"][i]), int(f0["prim"][i]), int(f1["status"][i]), int(f1["prim"][i])) for i in bad).most_common(12))
print("from-prim of mismatches:", Counter(int(h0["prim"][m][i]) for i in bad).most_common(12))
for i in bad[:5]:
    print("ray", i, "o", o[i], "d", wi[i], "fp64", f0["status"][i], f0["prim"][i], f0["chord"][i], f0["t"][i], "fp32", f1["status"][i], f1["prim"][i], f1["chord"][i], f1["t"][i])
    n = max(f0["chord"][i], f1["chord"][i]) + 1
    s0 = c0.geodesic(o[i:i+1], wi[i:i+1], n)[0]; s1 = c1.geodesic(o[i:i+1], wi[i:i+1], n)[0]
    for j in list(range(min(n, 3))) + list(range(max(3, n - 3), n)):
        print("   ", j, "fp64", s0[j], "\n       fp32", s1[j])
