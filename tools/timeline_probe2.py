"""Per-batch timeline (RTC_TIMELINE=1) of the teapot frame: rtc_render into a page-locked image (kernels done, copy done per
batch) next to the device-resident frame with the same batch schedule (kernels done)."""
import os, sys
os.environ["RTC_TIMELINE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracer_challenge_b200 as R
s = R.Session("teapot", 3840, 2160, device=0)
frame, ptr = R.pinned_array((2160, 3840, 3))
for i in range(3):
    print(f"-- host output, call {i}", file=sys.stderr, flush=True)
    _, st = s.render(R.Opts(), out=frame)
    print(f"   events {st.ms_kernels:.3f} ms total {st.ms_total:.3f}", file=sys.stderr, flush=True)
R.set_tuning(batches=14)
for i in range(3):
    print(f"-- device-resident, 14 batches, call {i}", file=sys.stderr, flush=True)
    _, _, st = s.render_device(R.Opts())
    print(f"   events {st.ms_kernels:.3f} ms", file=sys.stderr, flush=True)
