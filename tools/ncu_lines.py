"""Per-CUDA-source-line cost of one kernel from an .ncu-rep captured with --import-source on (read here, no GPU).
Usage: python tools/ncu_lines.py rep.ncu-rep KERNEL_REGEX [occurrence] [top]
Prints, per (file, line): warp-instructions executed, share, average active threads, stall samples."""
import csv, io, subprocess, sys, collections
rep, rx = sys.argv[1], sys.argv[2]
occ = int(sys.argv[3]) if len(sys.argv) > 3 else 0
top = int(sys.argv[4]) if len(sys.argv) > 4 else 45
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "-k", "regex:" + rx], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
# sections: "File Path", then "Function Name", then header row, then line rows (Line No non-empty) and sass rows (Line No empty)
per_kernel = collections.OrderedDict()
fname, func, hdr = None, None, None
seen_addr = set()
for r in rows:
    if not r: continue
    if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": func = r[1]; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) != len(hdr): continue
    if r[0] != "":  # a source line row: aggregated over its sass
        key = (func,)
        d = per_kernel.setdefault(func, collections.OrderedDict())
        col = {h: i for i, h in enumerate(hdr)}
        try:
            inst = int(r[col["Instructions Executed"]]); thr = int(r[col["Thread Instructions Executed"]]); smp = int(r[col["# Samples"]])
        except ValueError:
            continue
        k = (fname, int(r[0]))
        if k in d: continue  # (a line is listed once per file section; repeated sections repeat it)
        d[k] = (inst, thr, smp, r[1].strip()[:90])
names = list(per_kernel)
print("kernels:", [n.split("(")[0][-40:] for n in names], file=sys.stderr)
d = per_kernel[names[occ]]
tot_i = sum(v[0] for v in d.values()); tot_s = sum(v[2] for v in d.values()); tot_t = sum(v[1] for v in d.values())
print(names[occ][:110]); print(f"warp-inst {tot_i}  thread-inst {tot_t}  avg threads {tot_t / max(tot_i, 1):.1f}  samples {tot_s}")
for (f, ln), (inst, thr, smp, src) in sorted(d.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{f[:22]:22s}:{ln:4d} inst {inst:>10d} {100 * inst / tot_i:5.1f}%  thr {thr / max(inst, 1):5.1f}  smp {100 * smp / max(tot_s, 1):5.1f}%  {src}")
