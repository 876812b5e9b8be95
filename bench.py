#!/usr/bin/env python
"""Headline benchmark: BASELINE.json's metric (Mrays/s and frame ms).

  python bench.py --gpus N --steps K --warmup W [--config c4]   # this repo's CUDA path (one process per GPU)
  python bench.py --impl reference --steps K --warmup W         # the reference algorithm on the host cores (oracle port)

A "step" is one frame.  Default config c4 = BASELINE.json configs[3], the one north_star's target is quoted on:
chapter15 teapot (6 320 smooth triangles, 3 reflective planes, 2 lights) at 3840x2160, fuel 5.  --config c5 is
configs[4] (CSG + Simplex / fractal noise, 7680x4320 with 4x4 supersampling); c2 / c3h / c3b are the other full-size
configs.  At N > 1 the frame's bands of four pixel rows are dealt round-robin to the ranks (strong scaling: the frame is
fixed); no collective is on the data path.

  value  : whole-job Mrays/s with the scene resident in HBM and the frame left in HBM; rays = rays actually traced
           (primary + shadow + reflected + refracted, R_uniq of SURVEY.md §8(d)); time = CUDA events on the
           launching stream around each frame, max over ranks.
  e2e    : the same metric through the C ABI the reference would bind (rtc_scene_create + rtc_render into pinned
           host memory + rtc_scene_destroy = the body of Image::par_render): scene upload (host -> device) and
           frame download (device -> host) are inside the timed region.  At N > 1 every rank copies its bands into
           one shared page-locked host image (N PCIe links in parallel).
"""
import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "Mrays/s (primary+secondary+shadow)"

# BASELINE.json configs at their full sizes.  c4 (configs[3]) is the one north_star's target is quoted on: the default,
# and the line the driver records.  cpu_window: the bounded CPU sample of the same frame (x0, y0, w, h).
# f64_per_prim: flops of the scene's usual non-triangle intersection (plane through one matrix row: 11 + 2; sphere 32;
# mixed CSG scenes ~40); f64_per_pattern: flops of the pattern evaluation per (hit, light) (plain 0; a mixture ~32; jittered
# patterns ~3 Simplex evaluations = ~400), SURVEY.md §8(d).
CONFIGS = {
    "c2": dict(scene="glass_air_bubble", w=1920, h=1080, spp=(1, 1), fuel=5, lights=1, f64_per_prim=32.0, f64_per_pattern=32.0, cpu_window=(704, 412, 512, 256),
               workload="chapter11_glass_air_bubble (two concentric glass spheres over a checkered plane, 1 light) 1920x1080 fuel 5"),
    "c3h": dict(scene="hexagon", w=3840, h=2160, spp=(1, 1), fuel=5, lights=1, f64_per_prim=40.0, f64_per_pattern=400.0, cpu_window=(0, 0, 3840, 2160),
                workload="chapter14_hexagon (6 spheres + 6 open cylinders in nested groups, Simplex point-jitter stripes, 1 light) 3840x2160 fuel 5"),
    "c3b": dict(scene="benchmark", w=3840, h=2160, spp=(1, 1), fuel=5, lights=1, f64_per_prim=32.0, f64_per_pattern=32.0, cpu_window=(1408, 824, 1024, 512),
                workload="chapter14_benchmark (1000 spheres in 8 groups + backdrop plane, 1 light) 3840x2160 fuel 5"),
    "c4": dict(scene="teapot", w=3840, h=2160, spp=(1, 1), fuel=5, lights=2, f64_per_prim=13.0, f64_per_pattern=0.0, cpu_window=(1664, 952, 512, 256),
               workload="chapter15_teapot (teapot_high.obj, 6320 smooth triangles + 3 reflective planes, 2 lights) 3840x2160 fuel 5"),
    "c5": dict(scene="csg_noise", w=7680, h=4320, spp=(4, 4), fuel=5, lights=2, f64_per_prim=40.0, f64_per_pattern=400.0, cpu_window=(3712, 2096, 256, 128),
               workload="CSG (nested difference / intersection / union, glass lens) + Simplex / fractal-noise patterns, 2 lights, 7680x4320 with 4x4 supersampling (530.8 M primary rays), fuel 5"),
}
CFG = CONFIGS["c4"]  # set by main()

# Algorithmic cost model (DESIGN.md "Roofline"): flops per unit of traversal work, counted from the reference's
# arithmetic; the units themselves are counted by the kernels (rtc_stats).
F32_PER_NODE = 24.0      # two slab tests: 6 sub + 6 mul each (min/max/compare = 0)
NODE_BYTES = 64.0        # one BVH node visit: both child boxes + links, from L1/L2
TRI_TEST_BYTES = 96.0    # one triangle test: 16 B leaf header + 80 B (p1, e1, e2) record, from L1/L2
HIT_FETCH_BYTES = 480.0  # per surface hit: primitive + material + both matrices + triangle normals (384 + 96)
F64_PER_TRIANGLE = 46.0  # Moller-Trumbore, src/shape.rs:824-860
F64_PER_XFORM = 33.0     # affine Ray::transform: 18 mul + 15 add (src/ray.rs:14-19 on rows 0..2)
F64_PER_HIT = 130.0      # prepare_state + normal (+ schlick when transparent), SURVEY.md §8(d)
F64_PER_HIT_LIGHT = 95.0 + 18.0  # Phong term + shadow-ray setup, per light
RAY_RECORD = 64.0        # bytes of one queue entry: origin + direction (48) + throughput (8) + fuel/sample (8)
HIT_RECORD = 96.0        # bytes of one hit-list entry: over_point, normal, eye (72) + throughput (8) + sample/prim (8) + pad (8)
SHADOW_READ = 32.0       # a shadow thread reads the over_point of its hit (two 16-byte chunks)
ACCUM = 24.0             # f64 rgb accumulator of a sample


def read_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """Samples SM clock / throttle reasons of one GPU through NVML every 5 ms while the timed region runs."""

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.samples = []
        self.reasons = set()
        self.max_mhz = None
        self.stop_flag = threading.Event()
        self.thread = None
        self.error = None

    def _run(self):
        try:
            import pynvml as nv

            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.gpu)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            names = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown, "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                     "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown, "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
            while not self.stop_flag.is_set():
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
                time.sleep(0.005)
            nv.nvmlShutdown()
        except Exception as e:  # noqa: BLE001
            self.error = repr(e)

    def start(self):
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()

    def stop(self):
        self.stop_flag.set()
        if self.thread:
            self.thread.join(timeout=2)
        sm = sorted(self.samples)
        out = {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(sm)}
        if self.error:
            out["error"] = self.error
        return out


L2_NOTE = "GPU arm: 256 MiB device memset between timed frames (L2 flush); the per-level ray queues of a batch exceed L2 anyway"


def config_dict(config_id):
    """`config` of BOTH arms' lines (identical, so the driver can see they ran the same thing)."""
    return {"workload": CFG["workload"], "config_id": config_id, "cpu_sample": cpu_sample_text(), "l2": L2_NOTE}


def cpu_sample_text():
    x0, y0, w, h = CFG["cpu_window"]
    spp = CFG["spp"][0] * CFG["spp"][1]
    return (f"window {w}x{h} at ({x0},{y0}) of the {CFG['w']}x{CFG['h']} frame" + (f" x {spp} samples per pixel" if spp > 1 else "") +
            "; rays = World::intersect queries of the reference algorithm (R_ref, with its per-light duplication of secondary rays)")


def cpu_reference_step(threads=0):
    """One bounded sample of the workload with the reference algorithm (oracle port of the Rust code)."""
    from common import oracle_render  # tests/common.py: ctypes access to build/liboracle.so

    img, counts, sec = oracle_render(CFG["scene"], CFG["w"], CFG["h"], window=CFG["cpu_window"], fuel=CFG["fuel"], spp=CFG["spp"], threads=threads)
    return int(counts.sum()), sec


def cpu_baseline_entry(rays, sec, steps=1):
    x0, y0, w, h = CFG["cpu_window"]
    px = w * h
    ms = sec / steps * 1e3
    return {"value": rays / sec / 1e6, "unit": "Mrays/s", "cores": os.cpu_count() or 1, "kind": "port", "sample": cpu_sample_text() + f", {sec / steps:.1f} s per sample",
            "rays": "R_ref (the GPU arm's comparable figure is ref_equivalent_mrays_per_s; its `value` counts R_uniq)",
            "ms_per_sample_pixel": ms / px, "frame_ms_scaled_by_pixels": ms / px * CFG["w"] * CFG["h"]}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from raytracer_challenge_b200 import _build

    _build.build_oracle()
    for _ in range(args.warmup):
        cpu_reference_step()
    rays, secs = 0, 0.0
    for _ in range(args.steps):
        r, s = cpu_reference_step()
        rays += r
        secs += s
    value = rays / secs / 1e6
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_dict(args.config),
        "cpu_baseline": cpu_baseline_entry(rays, secs, args.steps),
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def run_cuda(args):
    # stdout carries exactly one line (the JSON): anything a library prints meanwhile (e.g. NCCL's version banner) goes to stderr
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        line = run_cuda_inner(args)
    finally:
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        os.close(saved_stdout)
    if line is not None:
        print(json.dumps(line), flush=True)
    return 0


def run_cuda_inner(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    import raytracer_challenge_b200 as R
    from raytracer_challenge_b200 import _build, sharding

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    if rank == 0:
        _build.build_rtc()
        _build.build_driver()
    torch.cuda.set_device(local)
    if world > 1:
        # one process per GPU: the ranks build their scenes at the same time and share the host cores
        R.set_tuning(build_threads=max((os.cpu_count() or world) // world, 1))
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # keep stdout to the one JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist.barrier()  # rank 0 has finished building

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(x, op):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=op)
        return float(t.item())

    MAX, SUM = (dist.ReduceOp.MAX, dist.ReduceOp.SUM) if world > 1 else (None, None)
    W, H, SPP, FUEL = CFG["w"], CFG["h"], CFG["spp"], CFG["fuel"]
    warmup = max(args.warmup, 3)
    opts = sharding.band_opts(W, rank, world, fuel=FUEL, spp=SPP)
    sess = R.Session(CFG["scene"], W, H, device=local)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    # ---- device-resident arm: `value` ----
    for _ in range(warmup):
        sess.render_device(opts)
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    t0 = time.perf_counter()
    ev_ms, rays, launches, rays_ref = 0.0, 0, 0, 0
    for _ in range(args.steps):
        flush.zero_()
        torch.cuda.synchronize()
        _, _, st = sess.render_device(opts)
        ev_ms += st.ms_kernels
        rays += st.rays
        rays_ref += st.rays_ref
        launches += st.kernel_launches
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop()
    ev_ms_max = reduce(ev_ms, MAX)
    rays_all = reduce(float(rays), SUM)
    rays_ref_all = reduce(float(rays_ref), SUM)
    launches_all = reduce(float(launches), SUM)
    ms_per_step = ev_ms_max / args.steps
    value = rays_all / args.steps / ms_per_step / 1e3  # Mrays/s

    # ---- one profiled frame (launches serialised, CUDA events around each): time per kernel class + work tallies ----
    popts = sharding.band_opts(W, rank, world, fuel=FUEL, spp=SPP, flags=R.FLAG_PROFILE)
    sess.render_device(popts)
    _, _, prof = sess.render_device(popts)

    # ---- end-to-end arm: scene upload + render + frame download through the C ABI, page-locked host output ----
    drv = R.load_driver()
    err = C.create_string_buffer(1024)
    st = R.Stats()
    frame = None
    pinned_ptr = None
    if world == 1:
        pinned, pinned_ptr = R.pinned_array((H, W, 3))
        out_ptr = pinned.ctypes.data
        host_image = pinned
    else:
        names = [None]
        if rank == 0:
            frame = sharding.SharedFrame(H, W, create=True)
            names[0] = frame.name
        dist.broadcast_object_list(names, src=0)
        if rank != 0:
            frame = sharding.SharedFrame(H, W, name=names[0], create=False)
        frame.pin()
        out_ptr = frame.ptr
        host_image = frame.array
    if rank == 0:
        host_image[:] = np.nan

    trace_steps = bool(os.environ.get("RTC_E2E_TRACE"))

    def e2e_step():
        ta = time.perf_counter()
        rc = drv.drv_e2e_step(sess.handle, local, C.byref(opts), out_ptr, C.byref(st), err, 1024)
        if rc != 0:
            raise R.RtcError(err.value.decode())
        tb = time.perf_counter()
        if world > 1:
            frame.frame_barrier(rank, world)  # the frame is complete when every rank's bands have landed
        if trace_steps:
            print(f"rank {rank}: step call {(tb - ta) * 1e3:.3f} ms, frame barrier {(time.perf_counter() - tb) * 1e3:.3f} ms", file=sys.stderr)
        return st.rays

    barrier()
    for _ in range(warmup):
        e2e_step()
    barrier()
    frame_ok = bool(np.isfinite(host_image).all()) if rank == 0 else True  # every pixel was written by some rank
    barrier()  # (rank 0's check must not count as the other ranks' first step)
    t0 = time.perf_counter()
    e2e_rays = 0
    for _ in range(args.steps):
        e2e_rays += e2e_step()
    t_loop = time.perf_counter()
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    if trace_steps:
        print(f"rank {rank}: e2e loop {(t_loop - t0) * 1e3:.3f} ms for {args.steps} steps, closing barrier {e2e_ms - (t_loop - t0) * 1e3:.3f} ms", file=sys.stderr)
    e2e_ms_max = reduce(e2e_ms, MAX)
    e2e_rays_all = reduce(float(e2e_rays), SUM)
    e2e_value = e2e_rays_all / e2e_ms_max / 1e3
    h2d = reduce(float(sess.build_stats["device_bytes"]), SUM)
    d2h = W * H * 3 * 8

    # ---- the real drop-in call: Image::par_render(camera, world) of the host API, one process driving all N GPUs ----
    # (flattening the World, rtc_scene_create_multi, rtc_render_multi into a pooled page-locked frame that becomes the
    # Image's pixels, rtc_scene_destroy).  With N > 1 ranks only rank 0 runs it, over devices 0..N-1, while the others wait.
    par = None
    barrier()
    if world > 1:
        frame.frame_barrier(rank, world)
    if rank == 0:
        probe = (C.c_double * 3)()
        phases = (C.c_double * 2)()
        pst = R.Stats()
        popts_plain = R.Opts(fuel=FUEL, spp=SPP)

        def par_step():
            rc = drv.drv_par_render_step(sess.handle, 0, world, C.byref(popts_plain), probe, C.byref(pst), phases, err, 1024)
            if rc != 0:
                raise R.RtcError(err.value.decode())
            if not all(np.isfinite(probe[k]) for k in range(3)):
                raise SystemExit("bench.py: Image::par_render returned a non-finite pixel")
            return pst.rays

        for _ in range(warmup):
            par_step()
        t0 = time.perf_counter()
        par_rays, flat_ms = 0, 0.0
        for _ in range(args.steps):
            par_rays += par_step()
            flat_ms += phases[0]
        par_ms = (time.perf_counter() - t0) * 1e3
        par = {"value": par_rays / par_ms / 1e3, "unit": "Mrays/s", "frame_ms": par_ms / args.steps, "flatten_ms": flat_ms / args.steps,
               "h2d_bytes_per_step": int(sess.build_stats["device_bytes"]) * world, "d2h_bytes_per_step": d2h,
               "path": f"rtc::Image::par_render(camera, world) (host/raytracer.hpp, the C++ mirror of the Rust call): FlatScene(world) + rtc_scene_create_multi + rtc_render_multi over {world} GPU(s) from ONE process "
                       "into a pooled page-locked frame that becomes Image.pixels + rtc_scene_destroy, per step; the World itself (OBJ parse, builders) is built once, outside"}
    if world > 1:
        # the other ranks wait HERE, on the host-side counters of the shared frame: an NCCL barrier would park a spinning kernel on
        # every GPU that rank 0 is rendering on
        frame.frame_barrier(rank, world, timeout_s=600.0)
    barrier()

    line = None
    if rank == 0:
        if not frame_ok:
            raise SystemExit("bench.py: the end-to-end frame holds pixels no rank wrote")
        line = {
            "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_dict(args.config),
            "run": {"sharding": f"bands of {sharding.TILE_H} pixel rows round-robin over {world} rank(s), scene replicated, no collective on the data path",
                    "scene": "device-resident arm: scene handle created once (host binned-SAH BVH), frames rendered from it; end-to-end arms: scene created and destroyed every step"},
            "frame_ms": ms_per_step, "wall_ms_per_step": wall_ms / args.steps,
            "rays_per_frame": rays_all / args.steps, "value_rays": "R_uniq: rays actually traced (secondary sub-trees deduplicated across lights)",
            "ref_equivalent_rays_per_frame": rays_ref_all / args.steps,
            "ref_equivalent_mrays_per_s": rays_ref_all / args.steps / ms_per_step / 1e3,
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": d2h, "frame_ms": e2e_ms_max / args.steps,
                    "ref_equivalent_mrays_per_s": e2e_value * rays_ref_all / max(rays_all, 1.0),
                    "path": "rtc_scene_create (build_flags = RTC_BUILD_FAST as Image::par_render passes: flattening on the host, BVH built on the device) + rtc_render (page-locked host f64 rgb, copies overlapped with tracing) + rtc_scene_destroy per step"
                            + ("" if world == 1 else "; every rank writes its bands into one shared host image, shared-memory barrier per frame")},
            "gpu_launches": int(launches_all),
        }
        # Headline e2e: at N = 1 the real call a bin makes, Image::par_render; the bare C-ABI sequence stays beside it.  At N > 1 the
        # headline stays the one-process-per-GPU form the driver launches (every rank renders its bands into one shared frame);
        # the one-process form (what the Rust binding would run) is reported beside it.
        if world == 1:
            line["e2e_c_abi"] = line["e2e"]
            line["e2e"] = par
        else:
            line["e2e_one_process"] = par
        line.update(rooflines(R, prof, world, args.config))
        if world == 1 and not args.no_cpu:
            _build.build_oracle()
            r, s = cpu_reference_step()
            line["cpu_baseline"] = cpu_baseline_entry(r, s)
        else:
            line["cpu_baseline"] = None
    if frame is not None:
        barrier()
        frame.close()
    if pinned_ptr is not None:
        R.free_pinned(pinned_ptr)
    sess.close()
    if world > 1:
        dist.destroy_process_group()
    return line


def kernel_source_hash():
    """Hash of the device code the traffic capture was taken on (profiles/roofline_traffic.json records it): a capture of
    other kernels than the ones being timed is stale and is not reported."""
    import hashlib

    h = hashlib.sha256()
    csrc = os.path.join(ROOT, "raytracer_challenge_b200", "csrc")
    for name in sorted(os.listdir(csrc)):
        if name.endswith((".cu", ".cuh", ".h")):
            with open(os.path.join(csrc, name), "rb") as f:
                h.update(name.encode() + b"\0" + f.read())
    return h.hexdigest()[:16]


def rooflines(R, prof, world, config_id):
    """Roofline figures of the dominant kernel class, from one profiled frame of rank 0 (its shard of the frame):
    algorithmic flops / bytes (cost model above x the work units the kernels counted) over the CUDA-event time of
    that class, against measured peaks (MEASURED_PEAKS.json for HBM; rtc_microbench for the CUDA-core issue rates
    and the L2 read bandwidth, which that file does not carry).  `roofline` is the bound with the largest fraction."""
    peak_hbm, peak_kind = read_peaks()
    L = float(CFG["lights"])
    hits = float(prof.surface_hits)
    secondary = float(prof.rays_reflected + prof.rays_refracted)
    levels = CFG["fuel"] + 1
    batches = max(int(prof.kernel_launches) // (2 * levels + 2), 1)  # per batch: trace_spawn x levels, trace_shadow, light x levels, resolve
    # traversal work split between the two traversal kernels in proportion to their rays (primary+secondary vs shadow)
    n_closest = float(prof.rays_primary) + secondary
    n_shadow = float(prof.rays_shadow)
    share = n_closest / max(n_closest + n_shadow, 1.0)
    nodes, tris, prims, xforms = float(prof.bvh_node_visits), float(prof.triangle_tests), float(prof.primitive_tests), float(prof.ray_transforms)
    f32_trav = F32_PER_NODE * nodes
    f64_trav = F64_PER_TRIANGLE * tris + CFG["f64_per_prim"] * prims + F64_PER_XFORM * xforms
    l2_trav = NODE_BYTES * nodes + TRI_TEST_BYTES * tris
    classes = {
        # trace_spawn: reads the level's ray records, writes the hit records and the children (the next level's rays);
        # also holds prepare_state + normal + the child rays
        "trace_spawn": {"ms": prof.ms_trace, "launches": batches * levels,
                        "hbm": RAY_RECORD * secondary + HIT_RECORD * hits + RAY_RECORD * secondary,
                        "f64": f64_trav * share + F64_PER_HIT * hits, "f32": f32_trav * share},
        # trace_shadow: per (hit, light) reads the hit's over_point, writes one byte; holds the shadow-ray setup
        "trace_shadow": {"ms": prof.ms_shadow, "launches": batches,
                         "hbm": L * hits * (SHADOW_READ + 1.0), "f64": f64_trav * (1.0 - share) + 18.0 * L * hits, "f32": f32_trav * (1.0 - share)},
        # light: reads the hit record, one shadow byte per light and its children's colours (24 B each), writes its own colour (24 B); the Phong terms
        "light": {"ms": prof.ms_shade, "launches": batches * levels,
                  "hbm": hits * (HIT_RECORD + L) + ACCUM * (hits + secondary), "f64": (95.0 + CFG["f64_per_pattern"]) * L * hits, "f32": 0.0},
    }
    # bytes the class moves through L1/L2 (SURVEY.md §8(d) BYTES(L1/L2)): its HBM records + the scene data it fetches
    classes["trace_spawn"]["l2"] = classes["trace_spawn"]["hbm"] + l2_trav * share + HIT_FETCH_BYTES * hits
    classes["trace_shadow"]["l2"] = classes["trace_shadow"]["hbm"] + l2_trav * (1.0 - share)
    classes["light"]["l2"] = classes["light"]["hbm"] + (32.0 + 96.0) * hits  # primitive + material record per hit
    classes = {k: v for k, v in classes.items() if v["ms"] > 0.0}
    dominant = max(classes, key=lambda k: classes[k]["ms"])
    d = classes[dominant]
    sec = max(d["ms"], 1e-9) * 1e-3
    n_launch = max(int(d["launches"]), 1)
    traffic, traffic_source, traffic_note = None, None, None
    tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tpath):  # per-launch DRAM bytes of the same kernel from the committed `ncu --set full` capture
        with open(tpath) as f:
            t = json.load(f)
        entry = t.get(config_id)
        if entry is None:
            traffic_note = f"no ncu capture for config {config_id} in profiles/roofline_traffic.json"
        elif entry.get("kernel_source_hash") != kernel_source_hash():
            traffic_note = "profiles/roofline_traffic.json was captured on other kernel sources than the ones timed here (stale): not reported"
        else:
            traffic = entry.get("per_launch_dram_bytes", {}).get(dominant)
            traffic_source = entry.get("source")
    try:
        mb = R.microbench()
    except Exception as e:  # noqa: BLE001
        mb = {"error": str(e)}
    bounds = {"hbm": {"achieved": d["hbm"] / sec / 1e9, "peak": peak_hbm, "unit": "GB/s", "peak_source": f"MEASURED_PEAKS.json hbm_gbs ({peak_kind})",
                      "what": "algorithmic HBM bytes of the class: ray queue records read and written, hit records, shadow bytes, accumulators"}}
    if "error" not in mb:
        bounds["l2"] = {"achieved": d["l2"] / sec / 1e9, "peak": mb["l2_read_gbs"], "unit": "GB/s", "peak_source": "rtc_microbench L2-resident read (measured in this run)",
                        "what": "SURVEY.md §8(d) BYTES(L1/L2): the HBM records plus BVH nodes, triangle records and per-hit scene fetches"}
        bounds["fp64"] = {"achieved": d["f64"] / sec / 1e12, "peak": mb["fp64_mul_add_tflops"], "unit": "TFLOP/s", "peak_source": "rtc_microbench (measured in this run)",
                          "what": "f64 CUDA-core issue, unfused mul / add (-fmad=false keeps f64 results bit-identical to the reference)"}
        bounds["fp32"] = {"achieved": d["f32"] / sec / 1e12, "peak": mb["fp32_mul_add_tflops"], "unit": "TFLOP/s", "peak_source": "rtc_microbench (measured in this run)",
                          "what": "fp32 CUDA-core issue, unfused mul / add (BVH slab tests)"}
    for b in bounds.values():
        b["frac"] = b["achieved"] / b["peak"]
        b["kernel"] = dominant
    top = max(bounds, key=lambda k: bounds[k]["frac"])
    out = {}
    out["roofline"] = dict(bounds[top])
    out["roofline"].update({
        "bound": top, "traffic": traffic, "traffic_unit": "bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum)", "traffic_source": traffic_source,
        "launches_per_frame": n_launch, "achieved_bytes_per_launch": d["hbm" if top == "hbm" else "l2"] / n_launch if top in ("hbm", "l2") else None,
        "algorithmic_hbm_bytes_per_launch": d["hbm"] / n_launch, "ms_per_launch": d["ms"] / n_launch,
        "ms_per_frame_in_this_kernel": d["ms"], "class_ms_source": "one serialised RTC_FLAG_PROFILE frame (one stream, CUDA events around every kernel class of every batch); "
                                                                   "the timed frames overlap several batch pipelines, so the class times sum to more than ms_per_step",
        "note": "the bound reported is the one with the largest fraction of {hbm, l2, fp64, fp32}; all four are listed under roofline_*.  The 2 MB scene is L1/L2-resident and the path "
                "is latency- and divergence-bound (profiles/README.md: issue-slot utilisation, active threads per instruction, stall reasons), so every fraction is far from 1",
        "scope": "rank 0's shard of the frame" if world > 1 else "whole frame",
        "per_class_ms": {k: v["ms"] for k, v in classes.items()},
    })
    if traffic_note:
        out["roofline"]["traffic_note"] = traffic_note
    for k, b in bounds.items():
        out["roofline_" + k] = b
    if "error" not in mb:
        out["measured_peaks"] = mb
    out["work_per_frame"] = {"surface_hits": prof.surface_hits, "bvh_node_visits": prof.bvh_node_visits, "triangle_tests": prof.triangle_tests,
                             "primitive_tests": prof.primitive_tests, "ray_transforms": prof.ray_transforms, "rays": prof.rays}
    return out


def main():
    global CFG
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=0, help="timed frames (default: 10; 3 for c5, whose frame takes over a second)")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--config", default="c4", choices=sorted(CONFIGS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    CFG = CONFIGS[args.config]
    if args.steps <= 0:
        args.steps = 3 if args.config == "c5" else 10
    if args.impl == "reference":
        return run_reference(args)
    return run_cuda(args)


if __name__ == "__main__":
    sys.exit(main())
