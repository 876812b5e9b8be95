"""Renders rank 0's bands of an N-rank run on one GPU a few times (for ncu launch lists).  Usage: shard_probe.py [world] [reps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracer_challenge_b200 as R
from raytracer_challenge_b200 import sharding
world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
s = R.Session("teapot", 3840, 2160, device=0)
for _ in range(reps):
    _, _, st = s.render_device(sharding.band_opts(3840, 0, world))
print(world, st.ms_kernels, st.kernel_launches, st.rays, "max nodes/ray", st.max_node_visits, "avg", st.bvh_node_visits / st.rays)
