"""More random scenes than the test-suite holds (fuzz:13 ...): BVH frame == brute-force frame (every item tested, boxes ignored)
== device-built-BVH frame, bit for bit, and the first ones against the oracle.  Stops after a time budget.
Usage: fuzz_sweep.py [first_seed] [last_seed] [seconds]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import raytracer_challenge_b200 as R
from common import compare_images, oracle_render
a, b = int(sys.argv[1]) if len(sys.argv) > 1 else 13, int(sys.argv[2]) if len(sys.argv) > 2 else 400
budget = float(sys.argv[3]) if len(sys.argv) > 3 else 30.0
t0 = time.time(); done = 0; bad = []; feats = {}; worst = 0.0; oracle_checked = 0
for seed in range(a, b + 1):
    if time.time() - t0 > budget: break
    name = f"fuzz:{seed}"
    try:
        with R.Session(name, 96, 54, device=0) as s:
            img, st = s.render(R.Opts())
            brute, _ = s.render(R.Opts(flags=1))
            feats[s.build_stats["features"]] = feats.get(s.build_stats["features"], 0) + 1
        if not np.array_equal(img, brute): bad.append((seed, "bvh != brute", float(np.abs(img - brute).max())))
        if seed % 4 == 0 and time.time() - t0 < budget * 0.8:
            ref, _, _ = oracle_render(name, 96, 54)
            rep = compare_images(img, ref); oracle_checked += 1
            worst = max(worst, rep["max_abs_f64"])
            if not rep["ok"] or rep["max_abs_f64"] >= 1e-9: bad.append((seed, "oracle", rep["max_abs_f64"]))
    except R.RtcError as e:
        bad.append((seed, "error", str(e)[:80]))
    done += 1
print(f"seeds {a}..{a + done - 1}: {done} scenes, kernel variants (features: scenes) {feats}, {oracle_checked} against the oracle (worst |diff| {worst:.2e}); failures: {bad}")
