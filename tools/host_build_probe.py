"""Host half of rtc_scene_create phase by phase (RTC_TRACE=1 lines of build_host_scene), best of N, no GPU needed.
Usage: host_build_probe.py [scene] [N]   (device-build staging path, what Image::par_render asks for)"""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
scene = sys.argv[1] if len(sys.argv) > 1 else "teapot"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 12
code = f"""
import sys; sys.path.insert(0, {ROOT!r})
import raytracer_challenge_b200 as R
R.set_tuning(bvh_builder=R.BVH_DEVICE)
for i in range({n}):
    R.Session({scene!r}, 64, 36, device=-1).close()
"""
out = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, RTC_TRACE="1"), capture_output=True, text=True).stderr
best = {}
for l in out.split("\n"):
    for k, v in re.findall(r"([a-z][a-z \-]+?) ([0-9.]+) ms", l.split(":", 1)[-1]):
        best[k.strip()] = min(best.get(k.strip(), 1e9), float(v))
print(scene, {k: round(v, 3) for k, v in best.items()}, "sum", round(sum(best.values()), 3))
