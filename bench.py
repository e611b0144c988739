#!/usr/bin/env python3
"""bench.py -- the headline measurement of BASELINE.json: Mrays/s (and ms/frame)
of the trace loop on scenes/scene5.json (glass dragon, path-traced GI) at
1920x1080, on N B200s of one node, beside the reference's CPU build.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--scene scene5] [--width 1920] [--height 1080] [--spp 16]

A "step" is one frame: one pass of the hot path (crt_render*, the replacement of
the reference's render(), src/raytracer/render.c:345-368) over all W*H pixels.
A "ray" is one get_closest_intersection() or is_light_blocked() call of the
reference; the count of one frame is a pure function of (scene,
options, seed) -- it is taken once, before timing, from the instrumented kernel
instance, and is identical to the reference's own count (tests/test_gpu_parity).

N > 1 (torchrun, one process per GPU): the scene is replicated and ALL ranks
drain one work queue (crt_frame_share, include/crt.h): frame and work counter
live on rank 0's GPU, mapped into the other processes by CUDA IPC over NVLink;
every resident warp of every GPU pulls the next 8-pixel unit with a
system-scope atomic and writes its pixels straight into rank 0's frame -- dynamic
load balance, no assembly step.  (--static-shards: the round-1 path, tiles dealt
round-robin + NCCL reduce.)  Total work is fixed, so scaling is "strong".

--impl reference times the UNMODIFIED reference (oracle/_ref/engine_ref, built
from /root/reference with its release flags and the thread-local RNG patch) on
the host cores, on a bounded sample of the same workload.
"""
import argparse
import ctypes as C
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

SAMPLE_W, SAMPLE_H = 320, 180  # CPU-baseline sample: the same view at 1/36 of the pixels (>= 5 rows per thread on 32 cores)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scene", default="scene5")
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--spp", type=int, default=16, help="-n: GI samples at the primary hit")
    ap.add_argument("--gi", default="path", choices=["path", "ambient"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--bvh", default="host", choices=["host", "device"],
                    help="where the acceleration structure is built (crt_scene_desc.build): binned SAH on the host cores (fewest box tests "
                         "per ray, the default) or Morton codes + radix tree on the GPU (milliseconds; a few percent slower frames)")
    ap.add_argument("--static-shards", action="store_true", help="N > 1: static tile shards + NCCL reduce instead of the shared work queue")
    ap.add_argument("--sample", default=f"{SAMPLE_W}x{SAMPLE_H}", help="frame size of the CPU-baseline sample")
    ap.add_argument("--ref-steps", type=int, default=1, help="--impl reference: timed runs of the sample (each is all host cores for tens of seconds)")
    ap.add_argument("--ref-warmup", action="store_true", help="--impl reference: one untimed run first")
    return ap.parse_args()


def workload_name(a):
    return f"scenes/{a.scene}.json {a.width}x{a.height} -g {a.gi} -n {a.spp} (defaults: -b 10 -a 0.01 -s phong -l sqr)"


def metric_name(a):
    what = {"scene5": "scene5 dragon", "scene6": "scene6 Menger sponge"}.get(a.scene, a.scene)
    res = {(1920, 1080): "1080p", (3840, 2160): "2160p"}.get((a.width, a.height), f"{a.width}x{a.height}")
    return f"Mrays/s ({what} {res} {'path-traced' if a.gi == 'path' else 'ambient'})"


def geometry_name(a, prims=None):
    mesh = {"scene5": "stand-in dragon.stl", "scene6": "stand-in menger_sponge.stl"}.get(a.scene)
    base = f"{mesh} (tools/gen_meshes.py)" if mesh else "the scene file's own geometry"
    return base if prims is None else f"{base}, {prims} bounded primitives"


def engine_flags(a):
    return ["-g", a.gi, "-n", str(a.spp)]


# ---- clocks ------------------------------------------------------------------

class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.samples = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.samples:
            if ts < t0 or ts > t1 + 0.2:
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
                power.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ---- reference arm -------------------------------------------------------------

def run_reference_sample(a, sample_w, sample_h):
    """Times oracle/_ref/engine_ref (reference main()/render(), -Ofast -fopenmp,
    -m max) on the sample and counts its rays with oracle/_ref/ref_engine (the
    same sources, instrumented).  Returns (rays, seconds, cores)."""
    from oracle import oracle
    if not os.path.exists(oracle.ENGINE_REF) or not os.path.exists(oracle.REF_ENGINE):
        raise RuntimeError("oracle/_ref is missing: run __graft_entry__.build() where /root/reference exists")
    scene = os.path.join("scenes", a.scene + ".json")
    # torchrun exports OMP_NUM_THREADS=1: the reference's -m max (system.c:54-74) must see the box's cores
    for k in ("OMP_NUM_THREADS", "OMP_PROC_BIND", "OMP_PLACES", "GOMP_CPU_AFFINITY"):
        os.environ.pop(k, None)  # also for oracle.ref_render below
    env = dict(os.environ)
    try:
        os.sched_setaffinity(0, range(os.cpu_count()))  # a launcher may also have pinned this process to one core
    except (AttributeError, OSError):
        pass
    with tempfile.TemporaryDirectory() as d:
        out = subprocess.run([oracle.ENGINE_REF, scene, os.path.join(d, "o.tif"), str(sample_w), str(sample_h), "-m", "max",
                              "-p", "real", "-f"] + engine_flags(a), cwd=ROOT, check=True, capture_output=True, text=True, env=env).stdout
        cores = int(re.search(r"Using (\d+) threads", out).group(1))
        t_start = float(re.search(r"\[(\d+\.\d+)\].*Commencing raytracing", out).group(1))
        t_end = float(re.search(r"\[(\d+\.\d+)\].*Saving image", out).group(1))
        key = (a.scene, sample_w, sample_h, tuple(engine_flags(a)))
        if key not in _RAY_COUNTS:  # a pure function of (scene, size, flags, seed): counted once
            _, _, cnt = oracle.ref_render(scene, sample_w, sample_h, engine_flags(a), seed=1)
            _RAY_COUNTS[key] = cnt["cast_ray"] + cnt["shadow"]
    return _RAY_COUNTS[key], t_end - t_start, cores


_RAY_COUNTS = {}


def reference_arm(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    sw, sh = (int(x) for x in a.sample.split("x"))
    # one run of the sample is 20 .. 60 s of all host cores: a single warm-up-free run unless asked otherwise
    warm = 1 if a.ref_warmup else 0
    for _ in range(warm):
        run_reference_sample(a, sw, sh)
    rays = secs = 0.0
    cores = 0
    steps = max(1, min(a.steps, a.ref_steps))
    for _ in range(steps):
        r, s, cores = run_reference_sample(a, sw, sh)
        rays += r
        secs += s
    mrays = rays / secs / 1e6
    full = a.width * a.height
    sample = (f"{a.scene} at {sw}x{sh} (same view, 1/{full // (sw * sh)} of the pixels, {sh} rows over {cores} threads), same flags; "
              f"reference render() timed by its own log, {steps} run(s); a full {a.width}x{a.height} frame would take "
              f"{secs / steps * full / (sw * sh):.0f} s at this rate")
    line = {
        "impl": "reference", "metric": metric_name(a), "value": mrays, "unit": "Mrays/s",
        "n_gpus": a.gpus, "steps": steps, "warmup": warm, "ms_per_step": secs / steps * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(a), "geometry": geometry_name(a), "sample": f"{sw}x{sh}",
                   "same_config": sw == a.width and sh == a.height, "host_cores": os.cpu_count(), "threads": cores},
        "cpu_baseline": {"value": mrays, "unit": "Mrays/s", "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": mrays, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ---- our arm -------------------------------------------------------------------

def ours(a):
    import numpy as np
    import torch
    import torch.distributed as dist

    import _pkg
    _pkg.load()
    from c_raytracer_b200 import _abi, device, host, multigpu

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available() or device.device_count() < 1:
        raise RuntimeError("bench.py: no CUDA device -- the trace loop has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # NCCL's banner must not share stdout with the JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    scene = host.Scene(os.path.join(ROOT, "scenes", a.scene + ".json"))
    t0 = time.time()
    dev = device.DeviceScene(scene, local, build=a.bvh)
    setup_s = time.time() - t0
    W, H = a.width, a.height
    pixels = W * H

    dynamic = world > 1 and not a.static_shards

    def make_opts(count=False):
        o = scene.options(W, H, *engine_flags(a))
        if not dynamic:
            multigpu.shard_options(o, rank, world)
        if count:
            o.flags |= _abi.CRT_RENDER_COUNT_STATS
        return o

    shared = multigpu.SharedFrame(dev, make_opts(), local) if dynamic else None

    d_raster = torch.empty((H, W, 3), dtype=torch.float32, device="cuda")
    d_z = torch.empty((H, W), dtype=torch.float32, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    stream = torch.cuda.current_stream()

    # ---- one instrumented frame: ray and test counts (untimed) --------------------
    st = _abi.Stats()
    if dynamic:
        shared.render(dev, make_opts(count=True))
    else:
        dev.render_device(make_opts(count=True), d_raster.data_ptr(), d_z.data_ptr(), stream.cuda_stream)
    dev.sync(st)
    if dynamic:
        shared.complete()
    counts = torch.tensor([st.rays_closest, st.rays_shadow, st.box_tests, *st.tri_tests, *st.sphere_tests, st.plane_tests],
                          dtype=torch.float64, device="cuda")
    flops_local = _abi.algorithmic_flops(st)
    local_rays = st.rays_closest + st.rays_shadow
    if world > 1:
        dist.all_reduce(counts)
    total_rays = float(counts[0] + counts[1])
    tot = _abi.Stats()
    tot.rays_closest, tot.rays_shadow, tot.box_tests = int(counts[0]), int(counts[1]), int(counts[2])
    for i in range(4):
        tot.tri_tests[i] = int(counts[3 + i])
    tot.sphere_tests[0], tot.sphere_tests[1], tot.plane_tests = int(counts[7]), int(counts[8]), int(counts[9])
    total_flops = _abi.algorithmic_flops(tot)

    opts = make_opts()
    asm_events = []
    ev_a0, ev_a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def step_device():
        """one frame, framebuffer left on rank 0's device"""
        flush.zero_()  # L2 flush between steps (outside the kernel's own events)
        if dynamic:
            shared.render(dev, opts)  # all ranks drain one work queue, pixels go straight to rank 0's frame
            ev_a0.record()
            shared.complete()         # one-element all-reduce: every rank's kernel has finished
            ev_a1.record()
            return
        dev.render_device(opts, d_raster.data_ptr(), d_z.data_ptr(), stream.cuda_stream)
        if world > 1:
            ev_a0.record()
            multigpu.assemble(d_raster, d_z, 0)  # NCCL reduce(sum) of the disjoint shards onto rank 0
            ev_a1.record()
            asm_events.append((ev_a0, ev_a1))

    # ---- device-resident timing ---------------------------------------------------
    for _ in range(a.warmup):
        step_device()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    barrier()
    kernel_ms = []
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    step_marks = [torch.cuda.Event(enable_timing=True) for _ in range(a.steps + 1)]
    t_wall0 = time.time()
    ev0.record()
    step_marks[0].record()
    for k in range(a.steps):
        step_device()
        step_marks[k + 1].record()
        kst = _abi.Stats()
        dev.sync(kst)  # per-frame kernel time from the library's own events on the launch stream
        kernel_ms.append(kst.kernel_ms)
    ev1.record()
    barrier()
    t_wall1 = time.time()
    clocks = sampler.stop(t_wall0, t_wall1) if rank == 0 else None
    elapsed = torch.tensor([ev0.elapsed_time(ev1), sum(kernel_ms) / len(kernel_ms)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(elapsed, op=dist.ReduceOp.MAX)
    total_ms, kernel_avg_ms = float(elapsed[0]), float(elapsed[1])
    # per rank and step: the render kernel alone, and the whole step (flush + kernel + waiting for the other ranks + assembly)
    torch.cuda.synchronize()
    mine = torch.tensor([kernel_ms, [step_marks[k].elapsed_time(step_marks[k + 1]) for k in range(a.steps)]], dtype=torch.float64, device="cuda")
    per_rank = [torch.zeros_like(mine) for _ in range(world)]
    if world > 1:
        dist.all_gather(per_rank, mine)
    else:
        per_rank = [mine]
    per_rank_ms = {"kernel": [[round(float(v), 2) for v in t[0]] for t in per_rank],
                   "step": [[round(float(v), 2) for v in t[1]] for t in per_rank]}
    assemble_ms = ev_a0.elapsed_time(ev_a1) if world > 1 else 0.0  # last frame's assembly, incl. waiting for the slowest rank
    assemble_alone_ms = 0.0
    if world > 1:  # the collective by itself, ranks aligned (untimed by the headline)
        barrier()
        ev_a0.record()
        for _ in range(5):
            if dynamic:
                shared.complete()
            else:
                multigpu.assemble(d_raster, d_z, 0)
        ev_a1.record()
        torch.cuda.synchronize()
        assemble_alone_ms = ev_a0.elapsed_time(ev_a1) / 5.0
    ms_per_step = total_ms / a.steps
    value = total_rays / (ms_per_step * 1e-3) / 1e6

    # ---- end to end: crt_render() with HOST buffers (the reference-facing call) ------
    h_raster = torch.empty((H, W, 3), dtype=torch.float32).pin_memory()
    h_z = torch.empty((H, W), dtype=torch.float32).pin_memory()
    if world > 1:
        acc_raster = torch.empty((H, W, 3), dtype=torch.float32).pin_memory() if rank == 0 else None

    def step_e2e():
        if world == 1:
            dev.render_into(opts, h_raster.numpy(), h_z.numpy())  # opts H2D as kernel params, frame D2H inside
        else:
            step_device()
            if rank == 0:
                h_raster.copy_(shared.raster if dynamic else d_raster, non_blocking=True)
                h_z.copy_(shared.z if dynamic else d_z, non_blocking=True)
            torch.cuda.synchronize()

    for _ in range(min(a.warmup, 2)):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        step_e2e()
    barrier()
    e2e_ms = torch.tensor([(time.perf_counter() - t0) * 1e3 / a.steps], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_value = total_rays / (float(e2e_ms[0]) * 1e-3) / 1e6
    frame_checksum = float(h_raster.double().sum()) if rank == 0 else 0.0

    # ---- roofline of the dominant kernel (render_kernel) -----------------------------
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    fp32_peak = device.measure_fp32_peak(local)  # FMA micro-kernel on this GPU, TFLOP/s
    achieved = flops_local / ((sum(kernel_ms) / len(kernel_ms)) * 1e-3) / 1e12
    info = dev.info
    # DRAM bytes per launch of this kernel from the committed ncu capture of this very workload (not measured live)
    traffic = None
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "dram_traffic.json")))
        # only the capture of THESE device sources on THIS workload counts (libcrt.so is not reproducible bit for bit)
        # ... or of sources recorded there as compiling to the same device code (host-only changes; the entry says how that was checked)
        same = [t.get("device_sources_sha256")] + [e.get("device_sources_sha256") for e in t.get("same_device_code", [])]
        if t["workload"] == workload_name(a) and world == t["n_gpus"] and not info.built_on_device and device.source_fingerprint() in same:
            traffic = t["dram_bytes_read"] + t["dram_bytes_write"]
    except (OSError, KeyError, ValueError):
        pass
    roofline = {
        "bound": "fp32", "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s", "frac": achieved / fp32_peak,
        "traffic": traffic,
        "kernel": "render_kernel<false, false>", "kernel_ms": sum(kernel_ms) / len(kernel_ms),
        "peak_source": "measured live: crt_measure_fp32_peak (FFMA micro-kernel); nominal 148*128*2*1.965 GHz = 74.4",
        "flops_per_ray": total_flops / total_rays, "box_tests_per_ray": tot.box_tests / total_rays,
        "note": "algorithmic FLOPs by SURVEY 8(d): 3/ray + 12/box test + tri 14/24/40/46 + sphere 17/20 + plane 12; "
                "HBM is not the bound (scene %.0f MB mostly L2-resident; traffic = DRAM bytes per launch from profiles/dram_traffic.json "
                "when it was captured from these device sources, else null; measured hbm peak %s GB/s)" % (info.device_bytes / 1e6, peaks.get("hbm_gbs")),
    }

    line = {
        "metric": metric_name(a), "value": value, "unit": "Mrays/s",
        "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {
            "workload": workload_name(a), "geometry": geometry_name(a, info.num_bounded),
            "rays_per_frame": total_rays, "rays_closest": tot.rays_closest, "rays_shadow": tot.rays_shadow,
            "l2": "flushed between steps (256 MiB memset)", "parallelism": (f"one dynamic work queue over {world} GPUs (crt_frame_share: 8-pixel units, system-scope atomics, "
                                                                          f"frame on rank 0 over CUDA IPC / NVLink), scene replicated" if dynamic else
                                                                          f"tile-chunk round-robin x{world}, scene replicated"),
            "bvh_nodes": info.num_nodes, "bvh_build_ms": info.build_ms, "bvh_built_on": "device" if info.built_on_device else "host", "scene_upload_ms": info.upload_ms,
            "scene_setup_s": setup_s, "seed": 1, "assemble_ms": assemble_ms, "assemble_alone_ms": assemble_alone_ms,
            "per_rank_ms": per_rank_ms,
        },
        "ms_per_frame": ms_per_step,
        "e2e": {"value": e2e_value, "unit": "Mrays/s", "ms_per_frame": float(e2e_ms[0]),
                "h2d_bytes_per_step": C.sizeof(_abi.RenderOpts), "d2h_bytes_per_step": 16 * pixels,
                "note": ("crt_render() with host buffers, scene resident (the reference's render() seam); " if world == 1 else
                         "crt_render_device_shared() on every rank + D2H of the finished frame from rank 0's GPU to pinned host memory; ") +
                        "scene build+upload is config.scene_setup_s, once per scene"},
        "gpu_launches": 2 * a.steps,
        "clocks": clocks,
        "roofline": roofline,
        "frame_checksum": frame_checksum,
    }

    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        sw, sh = (int(x) for x in a.sample.split("x"))
        # rays of the sample frame: counted by the instrumented kernel instance (equal to the reference's own
        # count, tests/test_gpu_parity.py) -- saves a second, instrumented, run of the reference
        so = scene.options(sw, sh, *engine_flags(a))
        so.flags |= _abi.CRT_RENDER_COUNT_STATS
        sst = _abi.Stats()
        dev.render(so, sst)
        _RAY_COUNTS[(a.scene, sw, sh, tuple(engine_flags(a)))] = sst.rays_closest + sst.rays_shadow
        rays, secs, cores = run_reference_sample(a, sw, sh)
        line["cpu_baseline"] = {
            "value": rays / secs / 1e6, "unit": "Mrays/s", "cores": cores, "kind": "reference",
            "host_cores": os.cpu_count(),
            "sample": f"{a.scene} at {sw}x{sh} (same view, 1/{pixels // (sw * sh)} of the pixels, {sh} rows over {cores} threads), same flags, "
                      f"{rays} rays in {secs:.2f} s; oracle/_ref/engine_ref = unmodified reference, -Ofast -fopenmp, -m max, thread-local RNG"}
    if rank == 0:
        print(json.dumps(line))
    if shared is not None:
        shared.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    a = parse_args()
    os.chdir(ROOT)  # mesh paths in scene files are cwd-relative
    if a.impl == "reference":
        return reference_arm(a)
    return ours(a)


if __name__ == "__main__":
    sys.exit(main())
