"""Rank 0's share of an N-rank frame on ONE GPU: device-resident kernel time (best of reps) and the wall time of the call.
Usage: shard_probe.py [scene:WxH] [worlds, e.g. 1,2,4,8] [reps]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracer_challenge_b200 as R
from raytracer_challenge_b200 import sharding
case = sys.argv[1] if len(sys.argv) > 1 else "teapot:3840x2160"
worlds = [int(x) for x in (sys.argv[2] if len(sys.argv) > 2 else "1,2,4,8").split(",")]
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 8
name, res = case.split(":")
w, h = map(int, res.split("x"))
s = R.Session(name, w, h, device=0)
for world in worlds:
    best, wall = 1e9, 1e9
    for _ in range(reps):
        t0 = time.perf_counter()
        _, _, st = s.render_device(sharding.band_opts(w, 0, world))
        wall = min(wall, (time.perf_counter() - t0) * 1e3)
        best = min(best, st.ms_kernels)
    print(f"world {world}: kernels {best:.3f} ms, call {wall:.3f} ms, launches {st.kernel_launches}, rays {st.rays}, Grays/s (x world) {st.rays / best / 1e6 * world:.1f}")
