"""Does a device -> host copy running on a copy engine slow the wavefront kernels down?  Device-resident teapot frames with and
without an unrelated 199 MB D2H copy loop on another stream (torch).
(What it measures is NOT a kernel slowdown: the frame's own tiny read-back queues behind the copy backlog on the D2H engine —
tools/copy_interference_probe4.py, profiles/README.md r02z.)"""
import os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import raytracer_challenge_b200 as R
s = R.Session("teapot", 3840, 2160, device=0)
n = 3840 * 2160 * 3
d = torch.zeros(n, dtype=torch.float64, device="cuda")
h = torch.empty(n, dtype=torch.float64).pin_memory()
side = torch.cuda.Stream()
def frames(k):
    best = 1e9
    for _ in range(k):
        best = min(best, s.render_device(R.Opts())[2].ms_kernels)
    return best
print("no copy: %.3f ms" % frames(8))
stop = False
def copier():
    with torch.cuda.stream(side):
        while not stop:
            h.copy_(d, non_blocking=True)
            side.synchronize()
t = threading.Thread(target=copier); t.start()
time.sleep(0.2)
print("with a concurrent D2H copy loop: %.3f ms" % frames(8))
stop = True; t.join()
for name, fn in (("H2D", lambda: d.copy_(h, non_blocking=True)),):
    stop = False
    def copier2():
        with torch.cuda.stream(side):
            while not stop:
                fn(); side.synchronize()
    t = threading.Thread(target=copier2); t.start(); time.sleep(0.2)
    print("with a concurrent %s copy loop: %.3f ms" % (name, frames(8)))
    stop = True; t.join()
