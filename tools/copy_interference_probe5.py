"""Device-resident teapot frames cut into 4 and into 14 batches, alone and next to an unrelated device -> host copy stream that
keeps the PCIe link busy (8 MB chunks, host sync after each, so nothing of the frame queues behind a backlog): does the cost of
copies in flight grow with the number of launches?"""
import os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import raytracer_challenge_b200 as R
s = R.Session("teapot", 3840, 2160, device=0)
CH = 1 << 20; NCH = 25
d = torch.zeros(CH * NCH, dtype=torch.float64, device="cuda")
h = torch.empty(CH * NCH, dtype=torch.float64).pin_memory()
side = torch.cuda.Stream()
def frames(k=10):
    return min(s.render_device(R.Opts())[2].ms_kernels for _ in range(k))
def with_copies(fn):
    stop = [False]
    def loop():
        with torch.cuda.stream(side):
            while not stop[0]:
                for k in range(NCH):
                    h[k * CH:(k + 1) * CH].copy_(d[k * CH:(k + 1) * CH], non_blocking=True)
                    side.synchronize()
    t = threading.Thread(target=loop); t.start(); time.sleep(0.15)
    r = fn()
    stop[0] = True; t.join()
    return r
for b in (4, 14, 28):
    R.set_tuning(batches=b)
    a = frames(); c = with_copies(frames); a2 = frames()
    print(f"{b} batches: alone {a:.3f} ms, next to the copy stream {c:.3f} ms, alone again {a2:.3f} ms", flush=True)
