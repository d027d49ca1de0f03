"""Multi-GPU inside the library (rrt_create_multi: one process, tiles dealt over the GPUs, scene replicated with ncclBroadcast, ONE
ncclReduce per frame) and through the C++ host (`host/pathtracer -G 0,1`, PathTracer::set_devices): the frame of G GPUs equals the
frame of one GPU bit for bit.  Reference counterpart: the N-worker tile pool of src/pathtracer.cpp:224-282.  Needs >= 2 GPUs
(gpurun --gpus 2); skipped on a single-GPU box."""
import importlib
import os
import subprocess

import numpy as np
import pytest

from util import ROOT, load_scene, sc

pytestmark = pytest.mark.gpu
EXE = os.path.join(ROOT, "host", "pathtracer")


@pytest.fixture(scope="module")
def api():
    a = importlib.import_module("relativistic-ray-tracer_b200.api")
    if a.lib().rrt_device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    return a


def ngpu(api):
    return min(api.lib().rrt_device_count(), 4)


@pytest.mark.parametrize("scene,mode", [("cbspheres_lambertian_480x360", sc.MODE_LITERAL), ("cbbunny_microfacet_480x360", sc.MODE_LITERAL),
                                        ("cbbunny_microfacet_480x360", sc.MODE_FLAT)])
def test_multi_gpu_frame_equals_single_gpu_frame(api, scene, mode):
    s = load_scene(scene)
    p = sc.make_params(ns_aa=40, max_ray_depth=4, ns_area_light=2, samples_per_batch=16, max_tolerance=0.2, seed=9)   # adaptive stop acts: 3 rounds
    one = api.Context(0); one.upload_scene(s); one.set_frame(480, 360); one.set_blackhole((0, 1, 0), 0.1, 0.1, mode)
    hdr, cnt, st = one.render(p)
    assert 0 < cnt.min() and cnt.min() < 40 and cnt.max() == 40
    devs = list(range(ngpu(api)))
    m = api.Context(devs)
    assert m.gpu_count() == len(devs)
    m.upload_scene(s); m.set_frame(480, 360); m.set_blackhole((0, 1, 0), 0.1, 0.1, mode)
    hm, cm, stm = m.render(p)
    assert np.array_equal(hdr, hm) and np.array_equal(cnt, cm)
    for k in ("paths", "rays", "geodesic_steps", "tri_tests"):
        assert stm[k] == st[k], k
    # a second frame (other seed) on the same context, and the parity hooks of a multi-GPU context (served by its first GPU)
    p2 = sc.make_params(ns_aa=4, max_ray_depth=3, ns_area_light=1, seed=10)
    assert np.array_equal(one.render(p2)[0], m.render(p2)[0])
    xy = np.random.default_rng(1).uniform(0, 1, (500, 2))
    o1, d1 = one.generate_rays(xy); o2, d2 = m.generate_rays(xy)
    assert np.array_equal(o1, o2) and np.array_equal(d1, d2)
    h1, _ = one.trace_rays(o1, d1); h2, _ = m.trace_rays(o1, d1)
    assert np.array_equal(h1["prim"], h2["prim"]) and np.array_equal(h1["hit_p"], h2["hit_p"])
    # a cell render and the device-resident form
    m.set_cell(100, 50, 64, 40); one.set_cell(100, 50, 64, 40)
    assert np.array_equal(one.render(p2)[0], m.render(p2)[0])


def test_multi_gpu_render_device_leaves_the_frame_on_the_first_gpu(api):
    """rrt_render_device on a multi-GPU context: d_rgbn / d_counts live on the first GPU and receive the reduced frame"""
    import torch
    s = load_scene("cbspheres_lambertian_96x72")
    p = sc.make_params(ns_aa=8, max_ray_depth=3, ns_area_light=2, seed=4)
    m = api.Context([0, 1]); m.upload_scene(s); m.set_frame(96, 72)
    hdr, cnt, _ = m.render(p)
    frame = torch.zeros((72, 96, 4), dtype=torch.float32, device="cuda:0")
    counts = torch.zeros((72, 96), dtype=torch.int32, device="cuda:0")
    torch.cuda.synchronize()
    st = m.render_device(p, frame.data_ptr(), counts.data_ptr())
    f = frame.cpu().numpy()
    assert np.array_equal(f[..., :3], hdr) and np.array_equal(f[..., 3], cnt.astype(np.float32)) and np.array_equal(counts.cpu().numpy(), cnt)
    assert st["paths"] == int(cnt.sum())


def test_multi_gpu_context_rejects_what_it_cannot_do(api):
    with pytest.raises(api.RrtError, match="listed twice"):
        api.Context([0, 0])
    m = api.Context([0, 1])
    with pytest.raises(api.RrtError, match="shards over its own GPUs"):
        m.set_tile_shard(8, 1, 2)
    m.set_tile_shard(16, 0, 1)
    with pytest.raises(api.RrtError, match="scene, camera and frame"):
        m.render(sc.make_params(ns_aa=1))


def test_cli_on_two_gpus_writes_the_same_png(api, tmp_path):
    if not os.path.exists(EXE):
        subprocess.run(["make", "-C", os.path.dirname(EXE)], check=True, capture_output=True)
    s = load_scene("cbspheres_lambertian_96x72")
    src = str(tmp_path / "s.rrtscene"); s.save(src)
    outs = []
    for g in ("0", "0,1"):
        out = str(tmp_path / ("o%d.png" % len(outs)))
        r = subprocess.run([EXE, "-s", "48", "-l", "2", "-m", "3", "-a", "16", "0.3", "-S", "5", "-G", g, "-f", out, src], capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
        outs.append((open(out, "rb").read(), open(out[:-4] + "_rate.png", "rb").read(), [l for l in r.stdout.splitlines() if "BVH traced" in l]))
    assert outs[0][0] == outs[1][0] and outs[0][1] == outs[1][1] and outs[0][2] == outs[1][2]
