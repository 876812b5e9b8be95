"""rtc_render of the teapot frame into a page-locked host image under a few tuning variants: wall ms (best of 8) and the
kernels' event time.  Usage: timeline_probe.py   (RTC_TIMELINE=1 prints when each batch's kernels / copy finished)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracer_challenge_b200 as R
s = R.Session("teapot", 3840, 2160, device=0)
frame, ptr = R.pinned_array((2160, 3840, 3))
variants = [dict(), dict(no_taper=1), dict(no_copy_stream=1), dict(no_stream_priorities=1), dict(no_graphs=1), dict(batches=14, pipes=4, no_taper=1, no_stream_priorities=1)]
for v in variants:
    R.set_tuning(**v) if v else R.set_tuning()
    best, bk = 1e9, 1e9
    for i in range(8):
        t = time.perf_counter()
        _, st = s.render(R.Opts(), out=frame)
        best = min(best, (time.perf_counter() - t) * 1e3)
        bk = min(bk, st.ms_kernels)
    dev = min(s.render_device(R.Opts())[2].ms_kernels for _ in range(5))
    print(f"{v}: host-output render {best:.3f} ms (events {bk:.3f}), device-resident with the same knobs {dev:.3f} ms", flush=True)
