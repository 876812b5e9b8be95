"""rtc_render of the teapot frame into a page-locked image: wall ms (best of 8) against the batch count / pipeline count."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracer_challenge_b200 as R
s = R.Session("teapot", 3840, 2160, device=0)
frame, ptr = R.pinned_array((2160, 3840, 3))
for pipes in (4, 3, 6):
    for batches in (6, 8, 10, 12, 14, 18, 24):
        for taper in (0, 1):
            R.set_tuning(batches=batches, pipes=pipes, no_taper=1 - taper)
            best = 1e9
            for i in range(8):
                t = time.perf_counter()
                s.render(R.Opts(), out=frame)
                best = min(best, (time.perf_counter() - t) * 1e3)
            print(f"pipes {pipes} batches {batches} taper {taper}: {best:.3f} ms", flush=True)
