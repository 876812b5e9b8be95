"""What about a device -> host copy slows the wavefront kernels (tools/copy_interference_probe.py: +35 % frame time; probe4 shows it is the read-back queueing, not the kernels)?  The same PCIe
traffic from (a) ONE 8 MB device region (always L2-resident) and (b) successive regions of a 200 MB buffer (every read a miss to
HBM), (c) at half the rate (a sleep between chunks), (d) device -> device copies on the copy engine (no PCIe)."""
import os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import raytracer_challenge_b200 as R
s = R.Session("teapot", 3840, 2160, device=0)
CH = 1 << 20  # doubles per chunk: 8 MB
NCH = 25
d = torch.zeros(CH * NCH, dtype=torch.float64, device="cuda")
d2 = torch.zeros(CH * NCH, dtype=torch.float64, device="cuda")
h = torch.empty(CH * NCH, dtype=torch.float64).pin_memory()
side = torch.cuda.Stream()
def frames(k=8):
    return min(s.render_device(R.Opts())[2].ms_kernels for _ in range(k))
def run(name, body):
    stop = [False]
    def loop():
        with torch.cuda.stream(side):
            while not stop[0]:
                body()
                side.synchronize()
    t = threading.Thread(target=loop); t.start(); time.sleep(0.2)
    ms = frames()
    stop[0] = True; t.join()
    print(f"{name}: {ms:.3f} ms", flush=True)
print(f"alone: {frames():.3f} ms", flush=True)
def one_region():
    for k in range(NCH): h[k * CH:(k + 1) * CH].copy_(d[:CH], non_blocking=True)
def all_regions():
    for k in range(NCH): h[k * CH:(k + 1) * CH].copy_(d[k * CH:(k + 1) * CH], non_blocking=True)
def whole():
    h.copy_(d, non_blocking=True)
def half_rate():
    for k in range(NCH):
        h[k * CH:(k + 1) * CH].copy_(d[k * CH:(k + 1) * CH], non_blocking=True)
        side.synchronize(); time.sleep(140e-6)
def d2d():
    d2.copy_(d, non_blocking=True)
run("D2H, 25 x 8 MB from ONE device region (L2-resident source)", one_region)
run("D2H, 25 x 8 MB from successive regions (HBM source)", all_regions)
run("D2H, one 200 MB copy", whole)
run("D2H at about half the rate (sleep between chunks)", half_rate)
run("D2D 200 MB copies (no PCIe)", d2d)
print(f"alone again: {frames():.3f} ms", flush=True)
