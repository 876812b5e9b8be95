"""Static code footprint of the render kernels: SASS instructions per kernel, and the source lines that own most of them
(nvdisasm -g on the cubin inside librtc.so; no GPU needed).  A line that owns thousands of instructions is usually a loop the
compiler unrolled at every inline site — code the instruction cache pays for whether it runs or not.
Usage: python tools/code_footprint.py [KERNEL_REGEX] [TOP]"""
import collections, os, re, subprocess, sys, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rx = re.compile(sys.argv[1] if len(sys.argv) > 1 else "trace_|light|resolve")
top = int(sys.argv[2]) if len(sys.argv) > 2 else 4
with tempfile.TemporaryDirectory() as tmp:
    subprocess.run(["cuobjdump", "-xelf", "render_kernels.sm_100a.cubin", os.path.join(ROOT, "raytracer_challenge_b200", "lib", "librtc.so")], cwd=tmp, capture_output=True)
    sass = subprocess.run(["nvdisasm", "-g", os.path.join(tmp, "render_kernels.sm_100a.cubin")], capture_output=True, text=True).stdout.split("\n")
name, cur, per = None, None, {}
for l in sass:
    t = l.lstrip()
    if t.startswith(".section"):
        m = re.match(r"\.section\s+\.text\.(\S+?),", t)
        name = m.group(1) if m else None
        cur = None
        continue
    if name is None:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
    elif re.match(r"\s*/\*[0-9a-f]{4,}\*/", l):
        per.setdefault(name, collections.Counter())[cur] += 1
for name, c in sorted(per.items(), key=lambda kv: -sum(kv[1].values())):
    demangled = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip().split("(")[0].replace("void rtcdev::", "")
    if not rx.search(demangled):
        continue
    tot = sum(c.values())
    print(f"{demangled:44s} {tot:6d} instructions ({tot * 16 // 1024:4d} KB)  top lines: " + ", ".join(f"{f}:{ln} {n}" for (f, ln), n in c.most_common(top)))
