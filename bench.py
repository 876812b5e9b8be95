#!/usr/bin/env python
"""Headline benchmark: BASELINE.json's metric (Mrays/s and frame ms) on the teapot config.

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one process per GPU)
  python bench.py --impl reference --steps K --warmup W    # the reference algorithm on the host cores (oracle port)

A "step" is one frame: chapter15 teapot (6 320 smooth triangles, 3 reflective planes, 2 lights) at 3840x2160,
fuel 5 (BASELINE.json configs[3], the config north_star's target is quoted on).  At N > 1 the frame's 8x4-pixel
tiles are dealt round-robin to the ranks (strong scaling: the frame is fixed); no collective is on the data path.

  value  : whole-job Mrays/s with the scene resident in HBM and the frame left in HBM; rays = rays actually traced
           (primary + shadow + reflected + refracted, R_uniq of SURVEY.md §8(d)); time = CUDA events on the
           launching stream around each frame, max over ranks.
  e2e    : the same metric through the C ABI the reference would bind (rtc_scene_create + rtc_render into pinned
           host memory + rtc_scene_destroy = the body of Image::par_render): scene upload and frame download are
           inside the timed region.
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

SCENE = "teapot"
WIDTH, HEIGHT, FUEL = 3840, 2160, 5
WORKLOAD = "chapter15_teapot (teapot_high.obj, 6320 smooth triangles + 3 reflective planes, 2 lights) 3840x2160 fuel 5"
CPU_WINDOW = (1664, 952, 512, 256)  # bounded CPU sample: a 512x256 window at the frame centre (on the teapot)
METRIC = "Mrays/s (primary+secondary+shadow)"


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


def cpu_reference_step(threads=0):
    """One bounded sample of the workload with the reference algorithm (oracle port of the Rust code)."""
    from common import oracle_render  # tests/common.py: ctypes access to build/liboracle.so

    img, counts, sec = oracle_render(SCENE, WIDTH, HEIGHT, window=CPU_WINDOW, fuel=FUEL, threads=threads)
    return int(counts.sum()), sec


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from raytracer_challenge_b200 import _build

    _build.build_oracle()
    cores = os.cpu_count() or 1
    for _ in range(args.warmup):
        cpu_reference_step()
    rays, secs = 0, 0.0
    for _ in range(args.steps):
        r, s = cpu_reference_step()
        rays += r
        secs += s
    value = rays / secs / 1e6
    px = CPU_WINDOW[2] * CPU_WINDOW[3]
    sample = f"window {CPU_WINDOW[2]}x{CPU_WINDOW[3]} at ({CPU_WINDOW[0]},{CPU_WINDOW[1]}) of the 3840x2160 frame per step; rays = World::intersect queries of the reference algorithm (R_ref, with its per-light duplication of secondary rays)"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": sample},
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cores, "kind": "port", "sample": sample,
                         "ms_per_sample_pixel": secs / args.steps * 1e3 / px, "frame_ms_scaled_by_pixels": secs / args.steps * 1e3 / px * WIDTH * HEIGHT},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def run_cuda(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    import raytracer_challenge_b200 as R
    from raytracer_challenge_b200 import _build

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    _build.build_rtc()
    _build.build_driver()
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

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

    opts = R.Opts(fuel=FUEL, shard=(rank, world))
    sess = R.Session(SCENE, WIDTH, HEIGHT, device=local)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    # ---- device-resident arm: `value` ----
    for _ in range(max(args.warmup, 3)):
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
    ev_ms_max = reduce(ev_ms, dist.ReduceOp.MAX if world > 1 else None)
    rays_all = reduce(float(rays), dist.ReduceOp.SUM if world > 1 else None)
    rays_ref_all = reduce(float(rays_ref), dist.ReduceOp.SUM if world > 1 else None)
    launches_all = reduce(float(launches), dist.ReduceOp.SUM if world > 1 else None)
    ms_per_step = ev_ms_max / args.steps
    value = rays_all / args.steps / ms_per_step / 1e3  # Mrays/s

    # ---- end-to-end arm: scene upload + render + frame download through the C ABI, pinned host output ----
    shape = sess.out_shape(opts)
    pinned, pinned_ptr = R.pinned_array(shape)
    h2d = e2e_bytes_h2d(sess)
    drv = R.load_driver()
    drv.drv_e2e_step.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int]
    err = C.create_string_buffer(1024)
    st = R.Stats()

    def e2e_step():
        rc = drv.drv_e2e_step(sess.handle, local, C.byref(opts), pinned.ctypes.data, C.byref(st), err, 1024)
        if rc != 0:
            raise R.RtcError(err.value.decode())
        return st.rays

    for _ in range(max(args.warmup, 3)):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    e2e_rays = 0
    for _ in range(args.steps):
        e2e_rays += e2e_step()
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    e2e_ms_max = reduce(e2e_ms, dist.ReduceOp.MAX if world > 1 else None)
    e2e_rays_all = reduce(float(e2e_rays), dist.ReduceOp.SUM if world > 1 else None)
    e2e_value = e2e_rays_all / e2e_ms_max / 1e3
    d2h = int(np.prod(shape)) * 8

    line = None
    if rank == 0:
        peak, peak_kind = read_peaks()
        # algorithmic HBM bytes of one frame for the wavefront kernels (DESIGN.md "roofline"): every secondary ray is
        # written once and read once as a 64-byte queue record; every surface hit adds 3 f64 to its sample (read +
        # write); every pixel is resolved once (24 B read + 24 B write).
        frame_rays = rays_all / args.steps
        alg_bytes = algorithmic_bytes(sess, frame_rays)
        achieved = alg_bytes / (ms_per_step * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "sharding": f"8x4-pixel tiles round-robin over {world} rank(s), scene replicated",
                       "l2": "256 MiB device memset between timed frames (L2 flush); ray queues (>500 MB per level) exceed L2 anyway"},
            "frame_ms": ms_per_step, "wall_ms_per_step": wall_ms / args.steps,
            "rays_per_frame": frame_rays, "ref_equivalent_rays_per_frame": rays_ref_all / args.steps,
            "ref_equivalent_mrays_per_s": rays_ref_all / args.steps / ms_per_step / 1e3,
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "frame_ms": e2e_ms_max / args.steps,
                    "path": "rtc_scene_create + rtc_render (pinned host f64 rgb) + rtc_scene_destroy per step"},
            "gpu_launches": int(launches_all),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                         "peak_source": f"MEASURED_PEAKS.json ({peak_kind})",
                         "note": "this path is FP64-issue / divergence bound, not HBM bound (scene is L2-resident); see roofline_fp64 and DESIGN.md"},
        }
        try:
            mb = R.microbench()
            line["roofline_fp64"] = fp64_roofline(sess, frame_rays, ms_per_step, mb)
        except Exception as e:  # noqa: BLE001
            line["roofline_fp64"] = {"error": str(e)}
        if world == 1 and not args.no_cpu:
            from raytracer_challenge_b200 import _build as b

            b.build_oracle()
            r, s = cpu_reference_step()
            px = CPU_WINDOW[2] * CPU_WINDOW[3]
            line["cpu_baseline"] = {"value": r / s / 1e6, "unit": "Mrays/s", "cores": os.cpu_count() or 1, "kind": "port",
                                    "sample": f"one {CPU_WINDOW[2]}x{CPU_WINDOW[3]} window at ({CPU_WINDOW[0]},{CPU_WINDOW[1]}) of the same frame, reference algorithm (R_ref rays), {s:.1f} s",
                                    "frame_ms_scaled_by_pixels": s * 1e3 / px * WIDTH * HEIGHT}
        else:
            line["cpu_baseline"] = None
    R.free_pinned(pinned_ptr)
    sess.close()
    if world > 1:
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line))
    return 0


def e2e_bytes_h2d(sess):
    b = sess.build_stats
    # device scene arrays uploaded by rtc_scene_create (device_scene.h record sizes)
    return (b["prims"] * 32 + b["xforms"] * 192 + b["bvh_nodes"] * 64 + b["bvh_items"] * 4 + 2 * 80 * max(b["prims"] - b["unbounded"], 0) + 4096)


def algorithmic_bytes(sess, frame_rays):
    px = WIDTH * HEIGHT
    # teapot frame: rays = primary + shadow + reflected; hits = shadow / 2 lights; queued (secondary) = reflected
    # solved from the stats of one frame in DESIGN.md; here: 64 B written + 64 B read per queued ray,
    # 48 B accumulate per hit, 48 B per pixel resolve.  Upper bound: treat every non-primary, non-shadow ray as queued.
    return 128.0 * max(frame_rays - px, 0) / 3.0 + 48.0 * frame_rays / 3.0 + 48.0 * px


def fp64_roofline(sess, frame_rays, ms_per_step, mb):
    # algorithmic f64 flops per ray for this scene: 3 plane tests with their ray transforms,
    # one mesh ray transform, ~12 Moller-Trumbore tests, amortised shading (prepare_state + Phong for 2 lights)
    flops_per_ray = 3 * (33 + 2) + 33 + 12 * 46 + (130 + 2 * 95) / 3.0
    achieved = flops_per_ray * frame_rays / (ms_per_step * 1e-3) / 1e12
    peak = mb["fp64_mul_add_tflops"]
    return {"bound": "fp64 issue (unfused mul/add, -fmad=false)", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
            "flops_per_ray_model": flops_per_ray, "measured": mb}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_cuda(args)


if __name__ == "__main__":
    sys.exit(main())
