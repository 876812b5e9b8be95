"""Per-kernel stall hot spots from `ncu --page source --csv` output (read here, no GPU).
Usage: ncu -i rep --page source --csv > src.csv ; python tools/ncu_hotspots.py src.csv [kernel-substring] [occurrence] [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2] if len(sys.argv) > 2 else ""
occ = int(sys.argv[3]) if len(sys.argv) > 3 else 0
topn = int(sys.argv[4]) if len(sys.argv) > 4 else 40
kernels = []  # (name, hdr, data)
i = 0
while i < len(rows):
    if rows[i] and rows[i][0] == "Kernel Name":
        name = rows[i][1]; hdr = rows[i + 1]; j = i + 2
        while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
            j += 1
        kernels.append((name, hdr, [r for r in rows[i + 2:j] if len(r) == len(hdr)]))
        i = j
    else:
        i += 1
sel = [k for k in kernels if want in k[0]]
print("kernels:", [(n.split("(")[0][-40:], len(d)) for n, h, d in kernels], file=sys.stderr)
name, hdr, data = sel[occ]
col = {h: k for k, h in enumerate(hdr)}
iS, iSrc, iI, iT = col["# Samples"], col["Source"], col["Instructions Executed"], col["Avg. Threads Executed"]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[iS]) for r in data)
print(name[:100]); print("samples", tot, "static instructions", len(data))
for s in stalls:
    v = sum(int(r[col[s]]) for r in data)
    if v > tot * 0.01: print(f"  {s}: {100*v/tot:.1f}%")
ex = [int(r[iI]) for r in data]
print("executed at least once:", sum(1 for e in ex if e > 0), "dyn warp-inst", sum(ex))
srt = sorted(ex, reverse=True); c = 0
for k, e in enumerate(srt):
    c += e
    if c > 0.9 * sum(ex): print("90% of executed warp-inst within", k, "static instructions"); break
top = sorted([(int(r[iS]), k) for k, r in enumerate(data)], reverse=True)[:topn]
for s, k in sorted(top, key=lambda x: x[1]):
    r = data[k]
    dom = max(stalls, key=lambda h: int(r[col[h]]))
    print(f"{k:5d} {s:6d} {100*s/tot:5.1f}% exec={r[iI]:>9} thr={r[iT]:>5} {dom[6:]:<14} {r[iSrc].strip()[:80]}")
