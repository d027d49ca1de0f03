#!/usr/bin/env python
"""bench.py — headline benchmark of the hot path (per-pixel Monte Carlo radiance estimation with Schwarzschild ray
bending) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c1|c3|c5|bunny_bh]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

One step = one full render of the workload.  Default workload = BASELINE.json configs[1] ("C2"):
CBspheres_lambertian.dae 480x360, -s 16 -l 4 -m 5, black hole on (the reference's default hole).  The scene is the
committed flattened copy of that .dae (tests/golden/*.rrtscene, written by the reference's own COLLADA loader).
At N GPUs the frame is tile-sharded (16x16 interleaved tiles), every rank renders 16*N samples per pixel of its
tiles (per-GPU work fixed => "scaling": "weak") and the per-rank float4 frames are combined with ONE NCCL reduce.

Prints ONE JSON line (rank 0).  `value` = Mpaths/s with everything resident in HBM; `e2e` = the same through the
public host-buffer API (scene upload + render + frame read-back inside the timed region).
--impl reference times the reference's own CPU path tracer (oracle/_ref/ref_bh, built from /root/reference by
oracle/build_ref.sh; falls back to the OpenMP oracle port when that binary is absent) on the host cores.
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
    "bunny_bh": ("cbbunny_microfacet_480x360", "CBbunny_microfacet_cu.dae", 1920, 1080, 16, 4, 5, "bh",
                 "C4 substitute (CBdragon.dae is not in the reference tree): CBbunny_microfacet_cu.dae 1920x1080 -s 16 -l 4 -m 5, hole on"),
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
def run_reference_binary(wl, spp, threads):
    """One run of the UNMODIFIED reference CPU path tracer (oracle/_ref/ref_bh or ref_flat).  Returns (paths, seconds, rays)."""
    fixture, dae, W, H, _, l, m, mode, _ = WORKLOADS[wl]
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_bh" if mode == "bh" else "ref_flat")
    scene = os.path.join(ROOT, "oracle", "_ref", "dae", dae or "-")
    template = os.path.join(ROOT, "oracle", "_ref", "dae", "CBspheres_lambertian.dae")
    synthetic = fixture == "synth:c5" and os.path.exists(template)
    if not os.path.exists(exe) or not (synthetic or (dae is not None and os.path.exists(scene))):
        return None
    with tempfile.TemporaryDirectory() as td:
        out = os.path.join(td, "o.png")
        cell = []
        if synthetic:
            # the procedural scene as a COLLADA file the reference's own loader reads (synth.write_c5_dae: vertices as float32),
            # and a bounded sample: the centre cell of the frame, 1/16 of its area (it holds the mesh: heavier than the average pixel)
            synth = importlib.import_module("relativistic-ray-tracer_b200.synth")
            scene = os.path.join(td, "c5.dae")
            synth.write_c5_dae(scene, open(template).read())
            cw, ch = W // 4, H // 4
            cell = ["-p", str((W - cw) // 2), str((H - ch) // 2), str(cw), str(ch)]
        cmd = [exe, "-t", str(threads), "-s", str(spp), "-l", str(l), "-m", str(m), "-r", str(W), str(H)] + cell + ["-f", out, scene]
        t0 = time.perf_counter()
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        wall = time.perf_counter() - t0
    txt = r.stdout
    ms = re.findall(r"100%! \(([0-9.]+)s\)", txt)          # reference src/pathtracer.cpp:636
    rays = re.findall(r"BVH traced (\d+) rays", txt)       # reference src/pathtracer.cpp:637
    secs = float(ms[-1]) if ms else wall
    npx = (W // 4) * (H // 4) if synthetic else W * H
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


def cpu_sample(wl, spp=1):
    cores = os.cpu_count() or 1
    res = run_reference_binary(wl, spp, cores)
    if res is not None:
        paths, secs, rays = res
        what = "centre cell (1/16 of the frame) of the same scene exported as .dae," if WORKLOADS[wl][0] == "synth:c5" else "same frame"
        return {"value": paths / secs / 1e6, "unit": "Mpaths/s", "cores": cores, "kind": "reference",
                "sample": "reference binary (oracle/_ref) -t %d, %s at -s %d: %d paths in %.2f s, %d rays" % (cores, what, spp, paths, secs, rays)}
    paths, secs, rays = run_oracle_port(wl, spp)
    return {"value": paths / secs / 1e6, "unit": "Mpaths/s", "cores": cores, "kind": "port",
            "sample": "oracle port (OpenMP, %d threads), same frame at %d spp: %d paths in %.2f s, %d rays" % (cores, spp, paths, secs, rays)}


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = args.workload
    times, paths = [], 0
    last = None
    for i in range(args.warmup + args.steps):
        last = cpu_sample(wl, 1)
        m = re.search(r"(\d+) paths in ([0-9.]+) s", last["sample"])
        if i >= args.warmup:
            paths += int(m.group(1)); times.append(float(m.group(2)))
    total = sum(times)
    v = paths / total / 1e6
    last["value"] = v
    print(json.dumps({"impl": "reference", "metric": "Mpaths/s", "value": v, "unit": "Mpaths/s", "n_gpus": args.gpus, "steps": args.steps,
                      "warmup": args.warmup, "ms_per_step": 1e3 * total / max(args.steps, 1), "higher_is_better": True, "scaling": "weak",
                      "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                      "config": {"workload": WORKLOADS[wl][8], "sample": "each step = the same frame at 1 spp on the host cores"},
                      "cpu_baseline": last, "e2e": {"value": v, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


# ------------------------------------------------------------------------------------------------ our arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
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
    spp_job = spp * world                       # weak scaling: per-GPU paths fixed
    s = workload_scene(args.workload)
    ctx = api.Context(local)
    ctx.set_option(api.OPT_BVH_TOPOLOGY, args.topology)
    ctx.set_option(api.OPT_LEAF_SIZE, args.leaf)
    ctx.set_option(api.OPT_TRACE_KERNEL, args.variant)
    ctx.set_option(api.OPT_CLEARANCE_GRID, args.grid)
    ctx.set_option(api.OPT_BVH_WIDTH, args.width)
    ctx.set_option(api.OPT_STEPPER, 1 if args.stepper == "fp32" else 0)
    ctx.set_option(api.OPT_BVH_BUILDER, {"auto": 0, "device": 1, "host": 2}[args.builder])
    ctx.upload_scene(s)
    ctx.set_blackhole((0.0, 1.0, 0.0), 0.1, 0.1, sc.MODE_LITERAL if mode == "bh" else sc.MODE_FLAT)
    ctx.set_frame(W, H)
    ctx.set_tile_shard(shard.DEFAULT_TILE, rank, world)
    # spp <= 32 keeps the reference's default adaptive batch (32) from ever triggering; above it the batch is the job's
    # spp so that every rank traces exactly W*H*spp/N paths (stated in config)
    spp_job, batch = shard.weak_scaling_params(spp, world)
    if args.batch:
        batch = args.batch
    params = sc.make_params(ns_aa=spp_job, max_ray_depth=m, ns_area_light=l, samples_per_batch=batch, seed=1)

    frame = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    def step_device():
        st = ctx.render_device(params, frame.data_ptr(), 0, 0, want_stats=True)
        shard.assemble(frame)
        return st

    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()          # early: NVML's first calls are slow; only samples inside the timed window are reported
    for _ in range(args.warmup):
        step_device()
    # ---- instrumented pass (counters on) for the algorithmic-byte and work figures; deterministic, so it describes
    # the timed passes exactly.  Timed passes run with counters off.
    st_count = step_device()
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
                      float(st_count["algorithmic_bytes"])], dtype=torch.float64, device="cuda")
    if world > 1:
        mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = t.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        total_ms, e2e_s, trace_s = float(mx[0]), float(mx[1]), float(mx[2])
        paths, steps_geo, rays, alg_bytes = float(sm[3]), float(sm[4]), float(sm[5]), float(sm[6])
    else:
        paths, steps_geo, rays, alg_bytes = [float(x) for x in t[3:7]]
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    secs = total_ms * 1e-3
    peak, peak_src = peaks()
    binfo = api.bvh_build_host(s, args.topology) if s.n_prims <= 200000 else None
    scene_bytes = (64 * binfo["flat_nodes"] + 80 * binfo["leaf_records"]) if binfo else 64 * (s.n_prims // 2) + 80 * s.n_prims
    # roofline of the dominant kernel (trace: geodesic-advance + BVH intersect): rank-0 algorithmic bytes / rank-0 trace time
    my_bytes = float(st_count["algorithmic_bytes"])
    my_trace_s = float(t[2]) / max(args.steps, 1)
    achieved = my_bytes / my_trace_s / 1e9 if my_trace_s > 0 else 0.0
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(args.workload)
    except (OSError, ValueError):
        pass
    line = {
        "metric": "Mpaths/s", "value": paths * args.steps / secs / 1e6, "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": desc, "frame": [W, H], "spp": spp_job, "ns_area_light": l, "max_ray_depth": m, "samples_per_batch": batch,
                   "parallelism": "tile%d" % world, "bvh_topology": args.topology, "bvh_leaf": args.leaf, "trace_variant": args.variant, "clearance_grid": args.grid, "bvh_width": args.width, "stepper": args.stepper, "bvh_builder": args.builder, "l2": "flushed between timed iterations (256 MiB write)",
                   "scaling_rule": "frame fixed, spp = %d x n_gpus, 16x16 interleaved tiles, one NCCL reduce per frame" % spp},
        "geodesic_steps_per_s": steps_geo * args.steps / secs, "rays_per_s": rays * args.steps / secs,
        "paths_per_step": paths, "geodesic_steps_per_step": steps_geo, "rays_per_step": rays,
        "e2e": {"value": paths * args.steps / e2e_s / 1e6, "unit": "Mpaths/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic["dram_bytes"] if traffic else None,
                     "traffic_of": (traffic["kernel"] + "; " + traffic["source"]) if traffic else None,
                     "kernel": "k_trace_closest + k_trace_shadow (geodesic advance + BVH intersect)", "peak_source": peak_src,
                     "algorithmic_bytes_per_step": my_bytes, "trace_ms_per_step": my_trace_s * 1e3,
                     "trace_launches_per_step": trace_launches / max(args.steps, 1),
                     "trace_share_of_step": trace_s / (render_s if render_s > 0 else 1.0),
                     "scene_bytes": scene_bytes,
                     "note": ("BVH nodes + leaf records = %.1f MB: %s" % (scene_bytes / 1e6,
                              "L1/L2 resident: the algorithmic bytes are served by the caches (compare `traffic`, the DRAM bytes of the same launch), so this fraction is not a DRAM utilisation and can exceed 1; the kernel is instruction-issue bound (ncu: 74 % of issue slots), see DESIGN.md section 6 and the fp64_step view"
                              if scene_bytes < 100e6 else "close to the 126 MB L2, yet served by it (ncu, profiles/traffic.json: L2 hit rate 95 %, 0.35 GB of DRAM traffic in the longest launch); the kernel is issue- / FP64-bound, see DESIGN.md section 6"))},
        "clocks": clk,
    }
    if steps_geo > rays and args.stepper == "fp64":   # hole on: the geodesic step is FP64 and its pipe, not HBM, is what the trace kernel runs against
        my_chords = float(st_count["geodesic_steps"])
        ceil = 48.5e9      # next_micro_ray + horizon_hit alone, saturated (tools/ilp_probe.cu on a B200)
        line["roofline"]["fp64_step"] = {"achieved": my_chords / my_trace_s / 1e9 if my_trace_s > 0 else 0.0, "peak": ceil / 1e9, "unit": "G chords/s",
                                         "frac": (my_chords / my_trace_s / ceil) if my_trace_s > 0 else 0.0,
                                         "peak_source": "measured: profiles/r01_i_fp64_step_probe.txt (174.5 FP64 warp instructions per chord)"}
    if not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_sample(args.workload, 1)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
