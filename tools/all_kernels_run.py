"""Small renders that touch every kernel (hole / flat, microfacet + glass, env map, hemisphere estimator, thin lens, split
shading on / off, 2- / 4- / 8-wide nodes, far chords, 2-way tile shard, the parity hooks): a quick everything-launches
check on a GPU box (compute-sanitizer is closed on this pool, so it is run plain)."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from util import load_scene, sc
api = importlib.import_module("relativistic-ray-tracer_b200.api")


def run(fixture, W, H, width=0, flat=False, hemi=False, env=False, split=2, shard=None, lens=False, spp=4):
    s = load_scene(fixture)
    c = api.Context(0)
    c.set_option(api.OPT_BVH_WIDTH, width); c.set_option(api.OPT_SHADE_SPLIT, split)
    if env:
        c.set_envmap(np.random.default_rng(3).random((16, 32, 3), dtype=np.float32))
    c.upload_scene(s)
    c.set_frame(W, H)
    c.set_blackhole((0, 1, 0), 0.1, 0.1, sc.MODE_FLAT if flat else sc.MODE_LITERAL)
    if shard:
        c.set_tile_shard(*shard)
    p = sc.make_params(ns_aa=spp, max_ray_depth=4, ns_area_light=2, seed=5, direct_hemisphere_sample=hemi, thin_lens=lens)
    hdr, cnt, st = c.render(p)
    print("%-30s %3dx%-3d width %d flat %d hemi %d env %d split %d lens %d shard %s: paths %d mean %.4f" %
          (fixture, W, H, width, flat, hemi, env, split, lens, shard, st["paths"], float(hdr.mean())), flush=True)
    ro = np.array([[0.0, 0.75, 4.8], [3e6, 2e6, -1e6]]); rd = np.array([[0.0, 0.0, -1.0], [-0.8017837, -0.5345225, 0.2672612]])
    rd /= np.linalg.norm(rd, axis=1, keepdims=True)
    c.trace_rays(ro, rd); c.trace_rays(ro, rd, any_hit=True); c.geodesic(ro, rd, 8)
    c.close()


run("cbspheres_lambertian_96x72", 96, 72)
run("cbspheres_lambertian_96x72", 96, 72, width=2, flat=True)
run("cbspheres_lambertian_96x72", 64, 48, width=4, env=True, split=1)
run("cbspheres_lambertian_96x72", 64, 48, hemi=True, shard=(16, 1, 2))
run("cbspheres_lambertian_96x72", 64, 48, lens=True, flat=True)
run("cbspheres_96x72", 64, 48, split=1)
run("cbspheres_microfacet_96x72", 64, 48)
run("cbgems_96x72", 64, 48, env=True)
run("cbbunny_microfacet_96x72", 64, 48, spp=2)
run("cbbunny_microfacet_96x72", 64, 48, width=8, flat=True, spp=2)
run("cbempty_96x72", 64, 48)
print("all_kernels_run ok")
