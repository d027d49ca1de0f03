"""The C++ host side end to end on the GPU: host/pathtracer (reference CLI flags) -> PathTracer -> C-ABI -> PNG,
against the same render made through the Python binding."""
import importlib
import os
import struct
import subprocess
import zlib

import numpy as np
import pytest

from util import ROOT, load_scene, sc

pytestmark = pytest.mark.gpu
EXE = os.path.join(ROOT, "host", "pathtracer")


def read_png(path):
    raw = open(path, "rb").read()
    pos, idat, w, h = 8, b"", 0, 0
    while pos < len(raw):
        n, tag = struct.unpack(">I4s", raw[pos:pos + 8])
        if tag == b"IHDR":
            w, h = struct.unpack(">II", raw[pos + 8:pos + 16])
        if tag == b"IDAT":
            idat += raw[pos + 8:pos + 8 + n]
        pos += 12 + n
    px = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + 4 * w)
    assert (px[:, 0] == 0).all()
    return px[:, 1:].reshape(h, w, 4)


@pytest.fixture(scope="module")
def exe():
    if not os.path.exists(EXE):
        subprocess.run(["make", "-C", os.path.dirname(EXE)], check=True, capture_output=True)
    return EXE


@pytest.mark.parametrize("flags,mode", [([], sc.MODE_LITERAL), (["-F"], sc.MODE_FLAT), (["-B", "0.1", "0.9", "0", "0.15", "0.2"], "alt")])
def test_cli_render_equals_python_binding(exe, tmp_path, flags, mode):
    api = importlib.import_module("relativistic-ray-tracer_b200.api")
    s = load_scene("cbspheres_lambertian_96x72")
    src = str(tmp_path / "s.rrtscene"); out = str(tmp_path / "o.png")
    s.save(src)
    r = subprocess.run([exe, "-s", "4", "-l", "2", "-m", "3", "-t", "8", "-S", "5", "-f", out] + flags + [src], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "[PathTracer] Job completed." in r.stdout and "BVH traced" in r.stdout
    c = api.Context(0); c.upload_scene(s); c.set_frame(96, 72)
    if mode == "alt":
        c.set_blackhole((0.1, 0.9, 0), 0.15, 0.2, sc.MODE_LITERAL)
    else:
        c.set_blackhole((0, 1, 0), 0.1, 0.1, mode)
    hdr, cnt, st = c.render(sc.make_params(ns_aa=4, max_ray_depth=3, ns_area_light=2, seed=5))
    rgba = c.tonemap(hdr).reshape(72, 96)[::-1]
    want = np.ascontiguousarray(rgba).view(np.uint8).reshape(72, 96, 4).copy()
    want[..., 3] = 255
    got = read_png(out)
    assert got.shape == want.shape and np.array_equal(got, want)
    assert ("traced %d rays" % st["rays"]) in r.stdout
    rate = read_png(out[:-4] + "_rate.png")
    assert rate.shape == (72, 96, 4) and (rate[..., 0] == 255).all() and (rate[..., 2] == 0).all()   # every pixel took all its samples: red


def test_cli_cell_render(exe, tmp_path):
    api = importlib.import_module("relativistic-ray-tracer_b200.api")
    s = load_scene("cbspheres_lambertian_96x72")
    src = str(tmp_path / "s.rrtscene"); out = str(tmp_path / "cell.png")
    s.save(src)
    r = subprocess.run([exe, "-s", "2", "-l", "1", "-m", "2", "-p", "20", "10", "32", "24", "-f", out, src], capture_output=True, text=True)
    assert r.returncode == 0 and "Cell job completed" in r.stdout, r.stdout + r.stderr
    c = api.Context(0); c.upload_scene(s); c.set_frame(96, 72)
    hdr, _, _ = c.render(sc.make_params(ns_aa=2, max_ray_depth=2, ns_area_light=1, seed=0))
    rgba = c.tonemap(hdr).reshape(72, 96)[10:34, 20:52][::-1]
    want = np.ascontiguousarray(rgba).view(np.uint8).reshape(24, 32, 4).copy(); want[..., 3] = 255
    assert np.array_equal(read_png(out), want)


def test_cli_loads_dae_when_present(exe, tmp_path):
    dae = os.path.join(ROOT, "oracle", "_ref", "dae", "CBspheres_lambertian.dae")
    if not os.path.exists(dae):
        pytest.skip("reference .dae files not on this box")
    out = str(tmp_path / "o.png")
    r = subprocess.run([exe, "-s", "1", "-l", "1", "-m", "1", "-r", "64", "48", "-f", out, dae], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    img = read_png(out)
    assert img.shape == (48, 64, 4) and img[..., :3].max() > 0


def test_cli_environment_map(exe, tmp_path):
    from util import golden
    api = importlib.import_module("relativistic-ray-tracer_b200.api")
    env = golden("kat_envmap")["env"]
    s = load_scene("cbspheres_lambertian_96x72")
    src = str(tmp_path / "s.rrtscene"); out = str(tmp_path / "o.png"); pfm = str(tmp_path / "env.pfm")
    s.save(src)
    with open(pfm, "wb") as f:
        f.write(b"PF\n%d %d\n-1.0\n" % (env.shape[1], env.shape[0]) + np.ascontiguousarray(env[::-1], np.float32).tobytes())
    r = subprocess.run([exe, "-s", "4", "-l", "2", "-m", "3", "-S", "5", "-e", pfm, "-f", out, src], capture_output=True, text=True)
    assert r.returncode == 0 and "Initializing environment light...done." in r.stdout, r.stdout + r.stderr
    c = api.Context(0); c.upload_scene(s); c.set_frame(96, 72); c.set_envmap(env)
    hdr, _, _ = c.render(sc.make_params(ns_aa=4, max_ray_depth=3, ns_area_light=2, seed=5))
    want = np.ascontiguousarray(c.tonemap(hdr).reshape(72, 96)[::-1]).view(np.uint8).reshape(72, 96, 4).copy(); want[..., 3] = 255
    assert np.array_equal(read_png(out), want)
    c.set_envmap(None)
    dark, _, _ = c.render(sc.make_params(ns_aa=4, max_ray_depth=3, ns_area_light=2, seed=5))
    assert hdr.mean() > dark.mean()      # the hole sends camera rays out of the box: they now see the sky


def test_cli_progress_snapshots_and_thin_lens(exe, tmp_path):
    """--progress writes one PNG per adaptive batch (the stand-in for the viewer's RENDERING frame); the last snapshot before the
    final batch differs from the final image, the final image equals the binding's; --thin-lens equals thin_lens=True."""
    api = importlib.import_module("relativistic-ray-tracer_b200.api")
    s = load_scene("cbspheres_lambertian_96x72")
    src = str(tmp_path / "s.rrtscene"); out = str(tmp_path / "o.png"); pre = str(tmp_path / "snap")
    s.save(src)
    r = subprocess.run([exe, "-s", "24", "-l", "1", "-m", "2", "-a", "8", "0", "-F", "--thin-lens", "--progress", pre, "-f", out, src], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    snaps = sorted(f for f in os.listdir(tmp_path) if f.startswith("snap_"))
    assert snaps == ["snap_000008.png", "snap_000016.png"]
    c = api.Context(0); c.upload_scene(s); c.set_frame(96, 72); c.set_blackhole((0, 1, 0), 0.1, 0.1, sc.MODE_FLAT)
    p = sc.make_params(ns_aa=24, max_ray_depth=2, ns_area_light=1, samples_per_batch=8, max_tolerance=0.0, seed=0, thin_lens=True)
    shots = []
    c.set_progress(lambda d, t, a: shots.append(c.snapshot()[0]) and False)
    hdr, _, _ = c.render(p)
    c.set_progress(None)
    def png_of(h):
        w = np.ascontiguousarray(c.tonemap(h).reshape(72, 96)[::-1]).view(np.uint8).reshape(72, 96, 4).copy(); w[..., 3] = 255
        return w
    assert np.array_equal(read_png(out), png_of(hdr))
    for f, h in zip(snaps, shots):
        assert np.array_equal(read_png(str(tmp_path / f)), png_of(h))
    assert not np.array_equal(read_png(str(tmp_path / snaps[-1])), read_png(out))


def test_cli_stepper_and_bvh_builder_flags(exe, tmp_path):
    """--stepper fp32 equals the binding with RRT_OPT_STEPPER = 1; --bvh-builder host / device write byte-identical PNGs."""
    api = importlib.import_module("relativistic-ray-tracer_b200.api")
    s = load_scene("cbbunny_microfacet_96x72")
    src = str(tmp_path / "s.rrtscene")
    s.save(src)
    outs = {}
    for name, flags in (("fp32", ["--stepper", "fp32"]), ("host", ["--bvh-builder", "host"]), ("device", ["--bvh-builder", "device"]), ("auto", [])):
        outs[name] = str(tmp_path / (name + ".png"))
        r = subprocess.run([exe, "-s", "4", "-l", "2", "-m", "3", "-S", "5", "-f", outs[name]] + flags + [src], capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
    assert np.array_equal(read_png(outs["host"]), read_png(outs["device"])) and np.array_equal(read_png(outs["host"]), read_png(outs["auto"]))
    c = api.Context(0); c.set_option(api.OPT_STEPPER, 1); c.upload_scene(s); c.set_frame(96, 72)
    hdr, _, _ = c.render(sc.make_params(ns_aa=4, max_ray_depth=3, ns_area_light=2, seed=5))
    w = np.ascontiguousarray(c.tonemap(hdr).reshape(72, 96)[::-1]).view(np.uint8).reshape(72, 96, 4).copy(); w[..., 3] = 255
    assert np.array_equal(read_png(outs["fp32"]), w)
    r = subprocess.run([exe, "--stepper", "fp16", src], capture_output=True, text=True)
    assert r.returncode != 0
