"""Emulates rank 0 of an N-rank END-TO-END frame on ONE GPU (its bands, scattered into a pinned full frame) and sweeps
the batch knobs: render ms of rtc_render (kernels + device->host copies), best of 6."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracer_challenge_b200 as R
from raytracer_challenge_b200 import sharding
W, H = 3840, 2160
s = R.Session("teapot", W, H, device=0)
frame, ptr = R.pinned_array((H, W, 3))
worlds = [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "1,2,4,8").split(",")]
for world in worlds:
    opts = sharding.band_opts(W, 0, world)
    flat = int(sys.argv[2]) if len(sys.argv) > 2 else 0  # 1: equal stream priorities
    for batches in (8, 10, 14, 20, 28):
        for pipes in (3, 4, 6):
            R.set_tuning(batches=batches, pipes=pipes, no_stream_priorities=flat)
            best = 1e9
            for rep in range(6):
                t = time.perf_counter()
                _, st = s.render(opts, out=frame)
                best = min(best, (time.perf_counter() - t) * 1e3)
            print(f"world {world} flat {flat} batches {batches:3d} pipes {pipes}: render {best:.3f} ms, launches {st.kernel_launches}", flush=True)
R.free_pinned(ptr)
