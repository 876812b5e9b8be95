"""rtc_render of the teapot frame into a page-locked image: device-side waits (default) against host-paced copies
(rtc_tuning.host_paced_copies) over a few batch counts; wall ms, best of 10."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import raytracer_challenge_b200 as R
s = R.Session("teapot", 3840, 2160, device=0)
frame, ptr = R.pinned_array((2160, 3840, 3))
ref, _ = s.render(R.Opts())
def best(**t):
    R.set_tuning(**t) if t else R.set_tuning()
    b = 1e9
    for i in range(10):
        t0 = time.perf_counter()
        _, st = s.render(R.Opts(), out=frame)
        b = min(b, (time.perf_counter() - t0) * 1e3)
    return b
print(f"default (device-side waits, graph replay): {best():.3f} ms", flush=True)
print(f"default, no graphs: {best(no_graphs=1):.3f} ms", flush=True)
for batches in (8, 10, 14, 18, 24):
    for taper in (1, 0):
        ms = best(host_paced_copies=1, batches=batches, no_taper=1 - taper)
        print(f"host-paced, {batches} batches, taper {taper}: {ms:.3f} ms  identical {bool(np.array_equal(frame, ref))}", flush=True)
print(f"host-paced, 14 batches, equal priorities: {best(host_paced_copies=1, no_stream_priorities=1):.3f} ms", flush=True)
print(f"default again: {best():.3f} ms", flush=True)
