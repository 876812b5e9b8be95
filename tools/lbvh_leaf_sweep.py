"""Device-built BVH (what Image::par_render asks for: RTC_BUILD_FAST): frame time against the leaf size, next to the host SAH tree."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracer_challenge_b200 as R
for name, w, h in (("teapot", 3840, 2160), ("teapot_tess", 3840, 2160), ("benchmark", 3840, 2160)):
    for label, t in (("host SAH", dict(bvh_builder=R.BVH_HOST)), ("device LBVH leaf 1", dict(bvh_builder=R.BVH_DEVICE, lbvh_leaf=1)), ("leaf 2", dict(bvh_builder=R.BVH_DEVICE, lbvh_leaf=2)),
                     ("leaf 3", dict(bvh_builder=R.BVH_DEVICE, lbvh_leaf=3)), ("leaf 4", dict(bvh_builder=R.BVH_DEVICE, lbvh_leaf=4))):
        R.set_tuning(**t)
        t0 = time.perf_counter()
        s = R.Session(name, w, h, device=0)
        t_open = (time.perf_counter() - t0) * 1e3
        ms = min(s.render_device(R.Opts())[2].ms_kernels for _ in range(6))
        print(f"{name}: {label}: frame {ms:.3f} ms, nodes {s.build_stats['bvh_nodes']}, depth {s.build_stats['bvh_depth']}", flush=True)
        s.close()
