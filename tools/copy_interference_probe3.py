"""Kernel slowdown against the RATE of a concurrent device -> host copy stream: 8 MB chunks with a host-side pause between
them (tools/copy_interference_probe2.py: at half rate the slowdown is gone)."""
import os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import raytracer_challenge_b200 as R
s = R.Session("teapot", 3840, 2160, device=0)
CHUNKS = {8: 1 << 20, 2: 1 << 18, 32: 1 << 22}
N = 25 << 20
d = torch.zeros(N, dtype=torch.float64, device="cuda")
h = torch.empty(N, dtype=torch.float64).pin_memory()
side = torch.cuda.Stream()
def frames(k=8):
    return min(s.render_device(R.Opts())[2].ms_kernels for _ in range(k))
print(f"alone: {frames():.3f} ms", flush=True)
def run(mb, pause_us, sync_each):
    ch = CHUNKS[mb]; nch = N // ch
    stop = [False]; moved = [0]; t_run = [0.0]
    def loop():
        t0 = time.perf_counter()
        with torch.cuda.stream(side):
            while not stop[0]:
                for k in range(nch):
                    h[k * ch:(k + 1) * ch].copy_(d[k * ch:(k + 1) * ch], non_blocking=True)
                    if sync_each:
                        side.synchronize()
                        if pause_us:
                            t = time.perf_counter()
                            while (time.perf_counter() - t) * 1e6 < pause_us: pass
                side.synchronize()
                moved[0] += N * 8
        t_run[0] = time.perf_counter() - t0
    t = threading.Thread(target=loop); t.start(); time.sleep(0.15)
    ms = frames()
    stop[0] = True; t.join()
    print(f"chunks of {mb} MB, pause {pause_us} us, sync per chunk {sync_each}: frame {ms:.3f} ms, copy stream {moved[0] / t_run[0] / 1e9:.1f} GB/s", flush=True)
run(8, 0, False)
for p in (0, 10, 20, 40, 70, 140):
    run(8, p, True)
run(2, 0, True)
run(32, 0, True)
run(32, 100, True)
print(f"alone again: {frames():.3f} ms", flush=True)
