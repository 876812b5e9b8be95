"""Is the kernels' slowdown next to queued device -> host copies real, or is it the frame's own small read-back (counters,
overflow flag) waiting in the D2H engine's queue behind them?  RTC_TIMELINE=1 prints when each batch's KERNELS finished."""
import os, sys, threading, time
os.environ["RTC_TIMELINE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import raytracer_challenge_b200 as R
s = R.Session("teapot", 3840, 2160, device=0)
CH = 1 << 20; NCH = 25
d = torch.zeros(CH * NCH, dtype=torch.float64, device="cuda")
h = torch.empty(CH * NCH, dtype=torch.float64).pin_memory()
side = torch.cuda.Stream()
def frames(tag, k=4):
    for i in range(k):
        print(f"-- {tag}, frame {i}", file=sys.stderr, flush=True)
        st = s.render_device(R.Opts())[2]
        print(f"   first-to-last event {st.ms_kernels:.3f} ms", file=sys.stderr, flush=True)
frames("alone")
def run(tag, sync_each):
    stop = [False]
    def loop():
        with torch.cuda.stream(side):
            while not stop[0]:
                for k in range(NCH):
                    h[k * CH:(k + 1) * CH].copy_(d[k * CH:(k + 1) * CH], non_blocking=True)
                    if sync_each: side.synchronize()
                side.synchronize()
    t = threading.Thread(target=loop); t.start(); time.sleep(0.15)
    frames(tag)
    stop[0] = True; t.join()
run("25 x 8 MB D2H queued back to back", False)
run("8 MB D2H chunks, host sync after each", True)
