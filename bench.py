#!/usr/bin/env python
"""bench.py — headline benchmark of the hot path (per-pixel Monte Carlo radiance estimation with Schwarzschild ray
bending) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c4|c2|c1|c3|c5|bunny_bh|bunny_box]
                    [--scaling strong|weak]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

One step = one full render of the workload.  Default workload = BASELINE.json configs[3] ("C4"), the configuration the
north_star's target is stated on: 1920x1080, -s 1024 -l 4 -m 5, black hole on, the reference's default adaptive sampling
(batch 32, tolerance 0.05, src/application.h:56-57).  CBdragon.dae is not in the reference tree
(/root/reference/.MISSING_LARGE_BLOBS), so the scene is the largest bundled Cornell-box scene,
CBbunny_microfacet_cu.dae ("CBdragon substitute"), as the committed flattened copy the reference's own COLLADA loader wrote.
At N GPUs the SAME frame at the SAME spp is tile-sharded (8x8 interleaved tiles; "scaling": "strong") and the per-rank
float4 frames are combined with ONE NCCL reduce; --scaling weak keeps round 1's rule (spp x N).

Prints ONE JSON line (rank 0).  `value` = Mpaths/s with everything resident in HBM (path = one traced camera sample, i.e. the
sum of the per-pixel sample counts after adaptive stopping); `e2e` = the same through the public host-buffer API (scene
upload + BVH build + render + frame read-back inside the timed region).  `roofline` is the FP64 pipe's: the trace kernel
(geodesic advance + BVH intersect, ~90 % of a step) is bound by FP64 instruction issue, its scene data is L1/L2 resident.
--impl reference times the reference's own CPU path tracer (oracle/_ref/ref_bh, built from /root/reference by
oracle/build_ref.sh) on the host cores, on a bounded sample of the same workload (see cpu_sample).
"""
import argparse
import importlib
import json
import os
import re
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np

WORKLOADS = {
    # name: (scene fixture, dae for the reference binary, W, H, spp, l, m, mode, description)
    "c2": ("cbspheres_lambertian_480x360", "CBspheres_lambertian.dae", 480, 360, 16, 4, 5, "bh",
           "C2: CBspheres_lambertian.dae 480x360 -s 16 -l 4 -m 5, Schwarzschild hole on (0,1,0) r=0.1 dtheta=0.1"),
    "c1": ("cbspheres_lambertian_480x360", "CBspheres_lambertian.dae", 480, 360, 16, 4, 5, "flat",
           "C1: CBspheres_lambertian.dae 480x360 -s 16 -l 4 -m 5, flat spacetime"),
    "c3": ("cbbunny_microfacet_480x360", "CBbunny_microfacet_cu.dae", 1280, 720, 256, 4, 5, "flat",
           "C3: CBbunny_microfacet_cu.dae 1280x720 -s 256 -l 4 -m 5, flat spacetime"),
    "c4": ("cbbunny_microfacet_480x360", "CBbunny_microfacet_cu.dae", 1920, 1080, 1024, 4, 5, "bh",
           "C4 (CBdragon substitute: CBdragon.dae is not in the reference tree): CBbunny_microfacet_cu.dae 1920x1080 -s 1024 -l 4 -m 5, "
           "Schwarzschild hole on (0,1,0) r=0.1 dtheta=0.1, adaptive sampling at the reference's defaults (batch 32, tolerance 0.05)"),
    "bunny_bh": ("cbbunny_microfacet_480x360", "CBbunny_microfacet_cu.dae", 1920, 1080, 16, 4, 5, "bh",
                 "C4 substitute (CBdragon.dae is not in the reference tree): CBbunny_microfacet_cu.dae 1920x1080 -s 16 -l 4 -m 5, hole on"),
    "bunny_box": ("cbbunny_microfacet_480x360", "CBbunny_microfacet_cu.dae", 480, 360, 64, 4, 5, "bh",
                  "CBbunny_microfacet_cu.dae 480x360 -s 64 -l 4 -m 5, hole on: the in-box pixels of the 1080p frame (the box covers the same "
                  "460x345 pixels at every -r, SURVEY App. C)"),
    "c5": ("synth:c5", None, 3840, 2160, 4, 4, 5, "bh",
           "C5 at reduced spp: procedural Cornell box + displaced UV sphere, 1,002,540 triangles (synth.py, seed 184), 3840x2160 -s 4 "
           "(the config names 4096 spp) -l 4 -m 5, hole on; BVH nodes + FP64 leaf records = 115 MB"),
}


def workload_scene(wl):
    """flattened scene of a workload at its frame size (fixture under tests/golden, or the procedural generator)"""
    from util import load_scene
    fixture, _, W, H = WORKLOADS[wl][:4]
    if fixture.startswith("synth:"):
        synth = importlib.import_module("relativistic-ray-tracer_b200.synth")
        s = synth.c5_scene(load_scene("cbspheres_lambertian_480x360"))
    else:
        s = load_scene(fixture)
    if (W, H) != (480, 360):
        s.set_screen_size(W, H)
    return s


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock and throttle reasons during the timed region: NVML (what nvidia-smi reads) polled every 10 ms from a
    thread, so that even a 25 ms timed region gets samples; falls back to `nvidia-smi -lms 100` when NVML cannot load."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows = []
        self.proc = None
        self.nvml = None
        self.index = index
        self._stop = threading.Event()

    def _nvml_handle(self):
        import pynvml
        pynvml.nvmlInit()
        try:
            import torch
            pr = torch.cuda.get_device_properties(self.index)
            bus = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
            return pynvml, pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        except Exception:
            return pynvml, pynvml.nvmlDeviceGetHandleByIndex(self.index)

    def start(self):
        try:
            self.nvml, self.handle = self._nvml_handle()
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _poll(self):
        n = self.nvml
        bits = [(getattr(n, "nvmlClocksEventReasonHwSlowdown", 0x8), "hw_slowdown"),
                (getattr(n, "nvmlClocksEventReasonHwThermalSlowdown", 0x40), "hw_thermal_slowdown"),
                (getattr(n, "nvmlClocksEventReasonSwThermalSlowdown", 0x20), "sw_thermal_slowdown"),
                (getattr(n, "nvmlClocksEventReasonSwPowerCap", 0x4), "sw_power_cap")]
        get_reasons = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or getattr(n, "nvmlDeviceGetCurrentClocksThrottleReasons")
        mx = n.nvmlDeviceGetMaxClockInfo(self.handle, n.NVML_CLOCK_SM)
        while not self._stop.is_set():
            try:
                sm = n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)
                r = get_reasons(self.handle)
                self.rows.append([str(sm), str(mx), "0"] + ["Active" if r & b else "Not Active" for b, _ in bits] + [time.perf_counter()])
            except Exception:
                pass
            self._stop.wait(0.01)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self, window=None):
        """window = (t0, t1) in time.perf_counter(): only the samples taken inside it count (NVML path; the sampler is started early
        so that NVML's first-call latency is not inside the timed region); fewer than 3 inside -> every sample, and the result says so."""
        if self.nvml:
            self._stop.set()
            self.t.join(timeout=2)
        elif not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        else:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows, scope = self.rows, "timed region"
        if window and self.nvml:
            inside = [r for r in rows if isinstance(r[-1], float) and window[0] <= r[-1] <= window[1]]
            rows, scope = (inside, "timed region") if len(inside) >= 3 else (rows, "whole run (warm-up + timed region)")
        for r in rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for k, nm in enumerate(names):
                if len(r) > 3 + k and r[3 + k].lower().startswith("active"):
                    reasons.add(nm)
        sm_load = sorted(sm)[len(sm) // 2:] if sm else []  # upper half = samples taken under load
        return {"sm_mhz": float(np.median(sm_load)) if sm_load else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvml" if self.nvml else "nvidia-smi", "scope": scope}


# ------------------------------------------------------------------------------------------------ CPU arms
# Workloads whose spp exceeds the adaptive batch: the reference's adaptive rule (src/part1_code.cpp:147-159) stops a pixel at the
# first batch boundary where 1.96 sigma / sqrt(n) <= tol * mu, so the frame's paths are far from W*H*spp and fall in two classes with
# very different costs.  Class A = pixels outside the central 480x360 cell: at every -r the Cornell box covers the same central
# pixels (the reference fixes screenDist at 800x600, SURVEY finding 0.1-4), outside it every camera ray escapes, mu = sigma = 0, and
# the pixel stops after exactly one batch of one-ray paths.  Class B = the central cell: the in-box pixels, ~15 rays per path.
# ADAPTIVE[wl] = paths of the whole frame and of class B, measured by the GPU arm (deterministic: seed 1; the GPU arm re-measures
# them every run and says so in config.class_mix if they moved) — the CPU arms need them to weight their two per-class rates.
ADAPTIVE = {
    "c4": {"batch": 32, "tol": 0.05, "cell": (720, 360, 480, 360), "paths_total": 222926816, "paths_cell": 162101216},
}


def run_reference_binary(wl, spp, threads, cell=None):
    """One run of the UNMODIFIED reference CPU path tracer (oracle/_ref/ref_bh or ref_flat), optionally on a -p cell
    (reference src/main.cpp:99-105).  Returns (paths, seconds, rays) or None when the binary / scene is not there."""
    fixture, dae, W, H, _, l, m, mode, _ = WORKLOADS[wl]
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_bh" if mode == "bh" else "ref_flat")
    scene = os.path.join(ROOT, "oracle", "_ref", "dae", dae or "-")
    template = os.path.join(ROOT, "oracle", "_ref", "dae", "CBspheres_lambertian.dae")
    synthetic = fixture == "synth:c5" and os.path.exists(template)
    if not os.path.exists(exe) or not (synthetic or (dae is not None and os.path.exists(scene))):
        return None
    with tempfile.TemporaryDirectory() as td:
        out = os.path.join(td, "o.png")
        if synthetic:
            # the procedural scene as a COLLADA file the reference's own loader reads (synth.write_c5_dae: vertices as float32)
            synth = importlib.import_module("relativistic-ray-tracer_b200.synth")
            scene = os.path.join(td, "c5.dae")
            synth.write_c5_dae(scene, open(template).read())
        cellarg = ["-p"] + [str(v) for v in cell] if cell else []
        # -a <spp+1> 0: no adaptive stop inside a CPU sample (it only acts above 32 spp anyway)
        cmd = [exe, "-t", str(threads), "-s", str(spp), "-l", str(l), "-m", str(m), "-r", str(W), str(H)] + cellarg + ["-f", out, scene]
        t0 = time.perf_counter()
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        wall = time.perf_counter() - t0
    txt = r.stdout
    ms = re.findall(r"100%! \(([0-9.]+)s\)", txt)          # reference src/pathtracer.cpp:636
    rays = re.findall(r"BVH traced (\d+) rays", txt)       # reference src/pathtracer.cpp:637
    secs = float(ms[-1]) if ms else wall
    npx = cell[2] * cell[3] if cell else W * H
    return npx * spp, secs, int(rays[-1]) if rays else 0


def run_oracle_port(wl, spp):
    from util import load_scene, sc
    from oracle import orc
    fixture, _, W, H, _, l, m, mode, _ = WORKLOADS[wl]
    s = workload_scene(wl)
    o = orc.Oracle(s)
    o.set_blackhole((0.0, 1.0, 0.0), 0.1, 0.1, sc.MODE_LITERAL if mode == "bh" else sc.MODE_FLAT)
    o.set_frame(W, H)
    t0 = time.perf_counter()
    _, cnt, st = o.render(sc.make_params(ns_aa=spp, max_ray_depth=m, ns_area_light=l, seed=1))
    return int(cnt.sum()), time.perf_counter() - t0, st["rays"]


_T1_CACHE = {}


def t1_versus_tn(wl, cores):
    """-t 1 beside -t <cores> (the reference's shared rand() lock, src/random_util.h:11-13, can make one thread the
    faster arm), once per process, on a small cell of the in-box region.  Returns (rate_t1, rate_tn, cell) in Mpaths/s or None."""
    if wl in _T1_CACHE:
        return _T1_CACHE[wl]
    fixture, dae, W, H = WORKLOADS[wl][:4]
    cw, ch = 96, 72
    cell = ((W - cw) // 2 - 2 * cw, (H - ch) // 2, cw, ch)     # left of the hole's image: walls, floor and mesh, no captured pixels
    if cell[0] < 0:
        cell = (max((W - cw) // 2, 0), max((H - ch) // 2, 0), min(cw, W), min(ch, H))
    a = run_reference_binary(wl, 1, 1, cell)
    b = run_reference_binary(wl, 1, cores, cell) if a else None
    _T1_CACHE[wl] = (a[0] / a[1] / 1e6, b[0] / b[1] / 1e6, cell) if a and b else None
    return _T1_CACHE[wl]


def cpu_sample(wl, spp=1, with_t1=True):
    """Bounded CPU sample of the workload on the host cores, as paths/s of the whole workload.
    * workloads without adaptive stopping (spp <= 32): every path costs the same on average -> the same frame at `spp` samples
      (C5: a strided set of 4 x 4 cells, 1/16 of the frame's area spread over the whole frame).
    * adaptive workloads (ADAPTIVE above): per-class rates r_A, r_B from the LEFT HALF of each class's cell at 1 spp (the scene is
      mirror-symmetric up to the mesh), combined with the workload's own class mix: rate = (P_A + P_B) / (P_A / r_A + P_B / r_B).
    The thread count is the faster of -t 1 and -t <cores> measured on a small cell (both rates are in `sample`)."""
    cores = os.cpu_count() or 1
    fixture, dae, W, H = WORKLOADS[wl][:4]
    mode = WORKLOADS[wl][7]
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_bh" if mode == "bh" else "ref_flat")
    have_scene = os.path.exists(os.path.join(ROOT, "oracle", "_ref", "dae", dae or "CBspheres_lambertian.dae"))
    if not (os.path.exists(exe) and have_scene):
        paths, secs, rays = run_oracle_port(wl, spp)
        return {"value": paths / secs / 1e6, "unit": "Mpaths/s", "cores": cores, "kind": "port",
                "sample": "oracle port (OpenMP, %d threads), same frame at %d spp: %d paths in %.2f s, %d rays" % (cores, spp, paths, secs, rays)}
    t1 = t1_versus_tn(wl, cores) if with_t1 else None
    threads = 1 if (t1 and t1[0] > t1[1]) else cores
    t1_txt = ("; -t 1 vs -t %d on cell %s at 1 spp: %.5f vs %.5f Mpaths/s -> -t %d used" % (cores, list(t1[2]), t1[0], t1[1], threads)) if t1 else ""
    ad = ADAPTIVE.get(wl)
    if ad and ad.get("paths_cell"):
        cx, cy, cw, ch = ad["cell"]
        pB = ad["paths_cell"]; pA = ad["paths_total"] - pB
        b = run_reference_binary(wl, 1, threads, (cx, cy, cw // 2, ch))             # class B: left half of the central cell
        a = run_reference_binary(wl, 1, threads, (0, cy, cw // 2, ch))              # class A: same size, frame's left edge
        rA, rB = a[0] / a[1], b[0] / b[1]
        v = (pA + pB) / (pA / rA + pB / rB) / 1e6
        return {"value": v, "unit": "Mpaths/s", "cores": threads, "kind": "reference", "seconds": a[1] + b[1],
                "sample": "reference binary (oracle/_ref) -t %d at 1 spp on two %dx%d cells of the same frame: in-box class (left half of the central "
                          "480x360 cell) %d paths / %d rays in %.2f s, background class (frame's left edge) %d paths in %.2f s; weighted by the workload's "
                          "class mix %d : %d paths%s" % (threads, cw // 2, ch, b[0], b[2], b[1], a[0], a[1], pB, pA, t1_txt)}
    if fixture == "synth:c5":
        # 16 cells of 1/256 of the frame each, on a 4 x 4 lattice over the whole frame (centre cell alone holds the mesh and is heavier
        # than the average pixel: ADVICE r01)
        cw, ch = W // 16, H // 16
        paths = rays = 0; secs = 0.0
        for j in range(4):
            for i in range(4):
                r = run_reference_binary(wl, spp, threads, ((2 * i + 1) * W // 8 - cw // 2, (2 * j + 1) * H // 8 - ch // 2, cw, ch))
                paths += r[0]; secs += r[1]; rays += r[2]
        what = "16 cells of %dx%d on a 4x4 lattice over the frame (1/16 of its area), scene exported as .dae," % (cw, ch)
    else:
        paths, secs, rays = run_reference_binary(wl, spp, threads)
        what = "same frame"
    return {"value": paths / secs / 1e6, "unit": "Mpaths/s", "cores": threads, "kind": "reference", "seconds": secs,
            "sample": "reference binary (oracle/_ref) -t %d, %s at -s %d: %d paths in %.2f s, %d rays%s" % (threads, what, spp, paths, secs, rays, t1_txt)}


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = args.workload
    vals, secs = [], []
    last = None
    for i in range(args.warmup + args.steps):
        last = cpu_sample(wl, 1)
        if i >= args.warmup:
            vals.append(last["value"]); secs.append(last.get("seconds", 0.0))
    # every step is the same sample: the run's rate is the harmonic mean of the steps' rates (total paths / total time)
    v = len(vals) / sum(1.0 / x for x in vals)
    last["value"] = v
    print(json.dumps({"impl": "reference", "metric": "Mpaths/s", "value": v, "unit": "Mpaths/s", "n_gpus": args.gpus, "steps": args.steps,
                      "warmup": args.warmup, "ms_per_step": 1e3 * sum(secs) / max(args.steps, 1), "higher_is_better": True, "scaling": args.scaling,
                      "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                      "config": {"workload": WORKLOADS[wl][8], "sample": "each step = one bounded CPU sample of the workload on the host cores (cpu_baseline.sample)"},
                      "cpu_baseline": last, "e2e": {"value": v, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


# ------------------------------------------------------------------------------------------------ our arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--workload", default="c4", choices=sorted(WORKLOADS))
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"], help="N > 1: strong = same frame and spp for every N (default); weak = spp x N")
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel")
    ap.add_argument("--topology", type=int, default=1)
    ap.add_argument("--leaf", type=int, default=2)
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--grid", type=int, default=0)
    ap.add_argument("--width", type=int, default=0, choices=[0, 2, 4, 8], help="BVH node width (0 = library default: 8 for small scenes, else 2)")
    ap.add_argument("--stepper", default="fp64", choices=["fp64", "fp32"],
                    help="geodesic stepper: fp64 = bit-identical to the reference (default, the headline); fp32 = RRT_OPT_STEPPER 1, within 1e-4")
    ap.add_argument("--builder", default="auto", choices=["auto", "host", "device"], help="BVH construction (RRT_OPT_BVH_BUILDER): auto = device LBVH for >= 4096 primitives, else host binned SAH")
    ap.add_argument("--batch", type=int, default=0, help="override samples_per_batch (the reference's -a batch; default rule: shard.weak_scaling_params)")
    ap.add_argument("--streams", type=int, default=0, help="RRT_OPT_STREAMS override (0 = library default)")
    ap.add_argument("--no-overlap", action="store_true", help="RRT_OPT_OVERLAP = 0")
    ap.add_argument("--sort-shadow", type=int, default=-1, help="RRT_OPT_SORT_SHADOW override (0 never, 1 chunks under 2 M paths, 2 always; -1 = library default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)

    import torch
    import torch.distributed as dist
    from util import load_scene, sc
    api = importlib.import_module("relativistic-ray-tracer_b200.api")
    shard = importlib.import_module("relativistic-ray-tracer_b200.shard")

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    fixture, _, W, H, spp, l, m, mode, desc = WORKLOADS[args.workload]
    if args.spp:
        spp = args.spp
    ad = ADAPTIVE.get(args.workload) if not args.spp else None
    if args.scaling == "weak":
        # round 1's rule: frame fixed, spp x N, one batch = the whole job (no adaptive stop) above 32 spp
        spp_job, batch = shard.weak_scaling_params(spp, world)
        rule = "weak: frame fixed, spp = %d x n_gpus, 8x8 interleaved tiles, one NCCL reduce per frame" % spp
    else:
        # strong: the SAME frame at the SAME spp and batch for every N (the reference hands a fixed frame's tiles to N workers,
        # src/pathtracer.cpp:243-255).  Adaptive workloads keep the reference's default batch; the others keep the whole job in one
        # batch above 32 spp so that every pixel takes exactly spp samples.
        spp_job = spp
        batch = ad["batch"] if ad else (32 if spp <= 32 else spp)
        rule = "strong: same frame, spp and batch for every N; 8x8 interleaved tiles, one NCCL reduce per frame"
    if args.batch:
        batch = args.batch
    s = workload_scene(args.workload)
    ctx = api.Context(local)
    ctx.set_option(api.OPT_BVH_TOPOLOGY, args.topology)
    ctx.set_option(api.OPT_LEAF_SIZE, args.leaf)
    ctx.set_option(api.OPT_TRACE_KERNEL, args.variant)
    ctx.set_option(api.OPT_CLEARANCE_GRID, args.grid)
    ctx.set_option(api.OPT_BVH_WIDTH, args.width)
    ctx.set_option(api.OPT_STEPPER, 1 if args.stepper == "fp32" else 0)
    ctx.set_option(api.OPT_BVH_BUILDER, {"auto": 0, "device": 1, "host": 2}[args.builder])
    if args.streams:
        ctx.set_option(api.OPT_STREAMS, args.streams)
    if args.sort_shadow >= 0:
        ctx.set_option(api.OPT_SORT_SHADOW, args.sort_shadow)
    if args.no_overlap:
        ctx.set_option(api.OPT_OVERLAP, 0)
    ctx.upload_scene(s)
    ctx.set_blackhole((0.0, 1.0, 0.0), 0.1, 0.1, sc.MODE_LITERAL if mode == "bh" else sc.MODE_FLAT)
    ctx.set_frame(W, H)
    ctx.set_tile_shard(shard.DEFAULT_TILE, rank, world)
    params = sc.make_params(ns_aa=spp_job, max_ray_depth=m, ns_area_light=l, samples_per_batch=batch,
                            max_tolerance=ad["tol"] if ad else 0.05, seed=1)

    frame = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    counts = torch.zeros((H, W), dtype=torch.int32, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    def step_device(want_counts=False):
        st = ctx.render_device(params, frame.data_ptr(), counts.data_ptr() if want_counts else 0, 0, want_stats=True)
        shard.assemble(frame)
        return st

    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()          # early: NVML's first calls are slow; only samples inside the timed window are reported
    for _ in range(args.warmup):
        step_device()
    # ---- instrumented pass (counters on) for the work figures; deterministic, so it describes the timed passes exactly.
    # Timed passes run with counters off.
    st_count = step_device(want_counts=True)
    cell_paths = 0.0
    if ad:
        cx, cy, cw, ch = ad["cell"]
        cell_paths = float(counts[cy:cy + ch, cx:cx + cw].sum().item())
    ctx.set_option(api.OPT_COUNTERS, 0)
    step_device()

    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    trace_s, render_s, launches, trace_launches = 0.0, 0.0, 0, 0
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t_win0 = time.perf_counter()
    for k in range(args.steps):
        flush.zero_()                    # L2 flush between timed iterations (outside the events)
        torch.cuda.synchronize()
        ev[k][0].record()
        st = step_device()
        ev[k][1].record()
        trace_s += st["trace_seconds"]; render_s += st["seconds"]; launches += st["kernel_launches"]; trace_launches += st["trace_launches"]
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    total_ms = sum(a.elapsed_time(b) for a, b in ev)
    clk = clocks.stop((t_win0, time.perf_counter())) if rank == 0 else None

    # ---- e2e: public host-buffer API; scene upload (pinned host arrays -> HBM, BVH build) + render + frame read-back
    hdr = np.empty((H, W, 3), np.float32); cnt = np.empty((H, W), np.int32)
    hdr_t = torch.from_numpy(hdr).pin_memory(); cnt_t = torch.from_numpy(cnt).pin_memory()
    hdr_p, cnt_p = hdr_t.numpy(), cnt_t.numpy()
    h2d = int(s.prim_kind.nbytes + s.prim_index.nbytes + s.tri_pos.nbytes + s.tri_nrm.nbytes + s.tri_mat.nbytes + s.sph.nbytes + s.sph_mat.nbytes +
              s.mats.nbytes + s.lights.nbytes + 18 * 8 + 48)
    d2h = int(hdr.nbytes + cnt.nbytes)
    host_frame = torch.empty((H, W, 4), dtype=torch.float32).pin_memory()
    # the scene's geometry arrays are page-locked in place (cudaHostRegister), as the contract asks of a step's inputs
    pinned = []
    rt = torch.cuda.cudart()
    for a in (s.tri_pos, s.tri_nrm, s.prim_index, s.prim_kind, s.tri_mat, s.sph):
        if a.nbytes >= 4096 and int(rt.cudaHostRegister(a.ctypes.data, a.nbytes, 0)) == 0:
            pinned.append(a)

    def step_e2e():
        ctx.upload_scene(s)
        if world == 1:
            ctx.render(params, hdr_p, cnt_p, want_stats=False)
        else:
            ctx.render_device(params, frame.data_ptr(), 0, 0, want_stats=False)
            shard.assemble(frame)
            if rank == 0:
                host_frame.copy_(frame, non_blocking=False)
    step_e2e()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    for a in pinned:
        rt.cudaHostUnregister(a.ctypes.data)

    t = torch.tensor([total_ms, e2e_s, trace_s, float(st_count["paths"]), float(st_count["geodesic_steps"]), float(st_count["rays"]),
                      float(st_count["algorithmic_bytes"]), float(st_count["tri_tests"]), float(st_count["sphere_tests"]), float(st_count["node_visits"]),
                      cell_paths, render_s], dtype=torch.float64, device="cuda")
    rank_ms = [render_s * 1e3 / max(args.steps, 1)]
    if world > 1:
        mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = t.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        allr = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allr, t)
        rank_ms = [float(x[11]) * 1e3 / max(args.steps, 1) for x in allr]
        total_ms, e2e_s = float(mx[0]), float(mx[1])
        trace_sum_s = float(sm[2])
        paths, steps_geo, rays, alg_bytes, tris, sphs, nodes, cell_paths = [float(x) for x in sm[3:11]]
    else:
        trace_sum_s = trace_s
        paths, steps_geo, rays, alg_bytes, tris, sphs, nodes, cell_paths = [float(x) for x in t[3:11]]
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    secs = total_ms * 1e-3
    hbm_peak, hbm_src = peaks()
    binfo = api.bvh_build_host(s, args.topology) if s.n_prims <= 200000 else None
    scene_bytes = (64 * binfo["flat_nodes"] + 80 * binfo["leaf_records"]) if binfo else 64 * (s.n_prims // 2) + 80 * s.n_prims
    # ---- roofline of the dominant kernel (trace: geodesic advance + BVH intersect), job-wide: algorithmic work of all ranks over the
    # summed trace time of all ranks per step (= one GPU's rate; at N = 1 simply rank 0's)
    trace_s_per_step = trace_sum_s / max(args.steps, 1) / world
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(args.workload)
    except (OSError, ValueError):
        pass
    alg_gbs = alg_bytes / world / trace_s_per_step / 1e9 if trace_s_per_step > 0 else 0.0
    algorithmic = {"bytes_per_step": alg_bytes, "rate": alg_gbs, "unit": "GB/s", "served_by": "L1/L2",
                   "hbm_peak": hbm_peak, "hbm_peak_source": hbm_src, "scene_bytes": scene_bytes,
                   "note": "SURVEY 8(d) bytes (64 B per node record, 80 B per triangle test, 32 B per sphere test, 160 B per ray): served by the caches "
                           "(scene = %.1f MB against a 126 MB L2; DRAM traffic of the longest launch in `traffic`), NOT a DRAM utilisation" % (scene_bytes / 1e6)}
    hole = steps_geo > rays
    if hole and args.stepper == "fp64":
        # FP64 instructions the path needs, per unit (thread instructions; /32 = warp instructions at full lanes):
        #   chord = next_micro_ray + horizon_hit = 174.5 (ncu on tools/ilp_probe.cu, profiles/r01_i_fp64_step_probe.txt)
        #   triangle test = 61, sphere test = 36 (Moller-Trumbore / quadratic run to completion; counted from the source: upper bounds,
        #   most tests leave at the first sign test) -- 2-6 % of the total
        fp64_warp_instr = (174.5 * steps_geo + 61.0 * tris + 36.0 * sphs) / 32.0
        peak = 578.7     # G FP64 warp-instr/s: plain independent DADD / DMUL stream on this part (DFMA: 533.2); tools/fp64_peak_probe.cu
        ach = fp64_warp_instr / world / trace_s_per_step / 1e9 if trace_s_per_step > 0 else 0.0
        roof = {"bound": "fp64_pipe", "achieved": ach, "peak": peak, "unit": "G FP64 warp-instr/s", "frac": ach / peak,
                "traffic": traffic["dram_bytes"] if traffic else None,
                "peak_source": "measured on B200: profiles/r02_a_fp64_peak_probe.txt (DADD / DMUL 578.7, DFMA 533.2 G warp-instr/s = 1.99 / 1.83 per clk per SM at 1965 MHz)",
                "per_unit": {"chord": 174.5, "triangle_test": 61, "sphere_test": 36},
                "stepper_alone": {"achieved": steps_geo / world / trace_s_per_step / 1e9 if trace_s_per_step > 0 else 0.0, "peak": 48.5, "unit": "G chords/s",
                                  "note": "next_micro_ray + horizon_hit alone, from registers, saturate at 48.5 G chords/s = 46 % of the pipe's plain-stream peak "
                                          "(tools/ilp_probe.cu): the reachable ceiling of this instruction mix (6 divisions + 4 square roots per chord)"}}
    else:
        roof = {"bound": "hbm", "achieved": alg_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": alg_gbs / hbm_peak,
                "traffic": traffic["dram_bytes"] if traffic else None, "peak_source": hbm_src,
                "note": "flat spacetime / FP32 stepper: no FP64 stepping; the bytes are the SURVEY 8(d) algorithmic bytes, served by L1/L2 (see `algorithmic`)"}
    roof.update({"traffic_of": (traffic["kernel"] + "; " + traffic["source"]) if traffic else None,
                 "kernel": "k_trace_closest + k_trace_shadow (geodesic advance + BVH intersect)",
                 "trace_ms_per_step": trace_s_per_step * 1e3, "trace_launches_per_step": trace_launches / max(args.steps, 1),
                 "trace_share_of_step": trace_s / (render_s if render_s > 0 else 1.0), "algorithmic": algorithmic})
    cfg = {"workload": desc, "frame": [W, H], "spp": spp_job, "ns_area_light": l, "max_ray_depth": m, "samples_per_batch": batch,
           "max_tolerance": ad["tol"] if ad else 0.05, "adaptive_stop": bool(ad) or spp_job > batch,
           "parallelism": "tile%d" % world, "bvh_topology": args.topology, "bvh_leaf": args.leaf, "trace_variant": args.variant, "clearance_grid": args.grid,
           "bvh_width": args.width, "stepper": args.stepper, "bvh_builder": args.builder, "l2": "flushed between timed iterations (256 MiB write)",
           "scaling_rule": rule}
    if ad:
        cfg["class_mix"] = {"cell": list(ad["cell"]), "paths_cell": cell_paths, "paths_total": paths,
                            "matches_table": bool(ad.get("paths_cell") == cell_paths and ad.get("paths_total") == paths)}
        ad = dict(ad, paths_cell=int(cell_paths), paths_total=int(paths))
        ADAPTIVE[args.workload] = ad
    line = {
        "metric": "Mpaths/s", "value": paths * args.steps / secs / 1e6, "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": cfg,
        "geodesic_steps_per_s": steps_geo * args.steps / secs, "rays_per_s": rays * args.steps / secs,
        "paths_per_step": paths, "geodesic_steps_per_step": steps_geo, "rays_per_step": rays, "node_visits_per_step": nodes,
        "rank_render_ms": {"min": min(rank_ms), "max": max(rank_ms), "all": [round(x, 3) for x in rank_ms]},
        "e2e": {"value": paths * args.steps / e2e_s / 1e6, "unit": "Mpaths/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": int(launches),
        "roofline": roof,
        "clocks": clk,
    }
    if not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_sample(args.workload, 1)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
