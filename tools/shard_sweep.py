"""Emulates the per-rank work of an N-GPU run on ONE GPU (rank 0's bands) and sweeps the batch knobs."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracer_challenge_b200 as R
from raytracer_challenge_b200 import sharding
W, H = 3840, 2160
s = R.Session("teapot", W, H, device=0)
for world in (1, 2, 4, 8):
    for batches in (2, 4, 8, 16, 32):
        for pipes in (2, 4):
            os.environ["RTC_BATCHES"] = str(batches); os.environ["RTC_PIPES"] = str(pipes)
            best = 1e9
            for rep in range(5):
                _, _, st = s.render_device(sharding.band_opts(W, 0, world))
                best = min(best, st.ms_kernels)
            print(f"world {world} batches {batches:3d} pipes {pipes}: {best:.3f} ms, launches {st.kernel_launches}, ideal {11.5/world:.2f}")
