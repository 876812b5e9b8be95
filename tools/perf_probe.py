"""Quick device-side timing of the named scenes (kernel ms, rays, Mrays/s).
Usage: perf_probe.py [--microbench] [--frames N] [scene:WxH[:SXxSY]]...
  --frames N   render exactly N device-resident frames per case and nothing else (the short form used under ncu)"""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracer_challenge_b200 as R

XFLAGS = int(os.environ.get('RTC_PROBE_FLAGS', '0'))  # extra rtc_opts.flags (experiments)

DEFAULT = ["teapot:3840x2160", "glass_air_bubble:1920x1080", "benchmark:3840x2160", "hexagon:3840x2160", "csg_noise:1920x1080"]
cases = sys.argv[1:]
frames = 0
if "--microbench" in cases:
    cases.remove("--microbench")
    print(json.dumps(R.microbench()))
if "--frames" in cases:
    k = cases.index("--frames")
    frames = int(cases[k + 1])
    del cases[k:k + 2]
cases = cases or DEFAULT
for c in cases:
    name, res, *rest = c.split(":")
    w, h = map(int, res.split("x"))
    spp = tuple(map(int, rest[0].split("x"))) if rest else (1, 1)
    t0 = time.time()
    s = R.Session(name, w, h, device=0)
    t_open = time.time() - t0
    best = None
    for i in range(frames or 4):
        _, n, st = s.render_device(R.Opts(spp=spp, flags=XFLAGS))
        if best is None or st.ms_kernels < best.ms_kernels:
            best = st
    if frames:
        print(json.dumps({"scene": name, "res": res, "spp": list(spp), "ms_kernels": round(best.ms_kernels, 3), "rays": best.rays, "launches": best.kernel_launches}))
        s.close()
        continue
    _, _, pr = s.render_device(R.Opts(spp=spp, flags=R.FLAG_PROFILE | XFLAGS))
    _, _, pr = s.render_device(R.Opts(spp=spp, flags=R.FLAG_PROFILE | XFLAGS))
    out, st2 = s.render(R.Opts(spp=spp, flags=XFLAGS))
    print(json.dumps({"scene": name, "res": res, "spp": list(spp), "open_s": round(t_open, 3), "ms_kernels": round(best.ms_kernels, 3), "rays": best.rays,
                      "mrays_s": round(best.rays / best.ms_kernels / 1e3, 1), "rays_ref": best.rays_ref, "launches": best.kernel_launches,
                      "serial_ms": {"trace": round(pr.ms_trace, 3), "shadow": round(pr.ms_shadow, 3), "shade": round(pr.ms_shade, 3)}, "hits": best.surface_hits,
                      "work": {"nodes": pr.bvh_node_visits, "tris": pr.triangle_tests, "prims": pr.primitive_tests, "xforms": pr.ray_transforms},
                      "e2e_ms_total": round(st2.ms_total, 2), "build": s.build_stats}), flush=True)
    s.close()
