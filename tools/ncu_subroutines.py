"""Share of a kernel's warp-instructions spent in out-of-line subroutines (f64 division / sqrt / pow special-case paths) and the
call sites that reach them, from the SASS source page of an ncu capture (no GPU needed).
Usage: python tools/ncu_subroutines.py ncu_source_TAG.csv[.gz]   (the file tools/gpu.sh ncu writes: `ncu -i REP --page source --csv`)"""
import csv, gzip, io, sys

path = sys.argv[1]
raw = gzip.open(path, "rt").read() if path.endswith(".gz") else open(path).read()
rows = list(csv.reader(io.StringIO(raw)))
secs = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"] + [len(rows)]
seen = set()
for a, b in zip(secs[:-1], secs[1:]):
    name = rows[a][1].split("(rtcdev")[0].replace("void rtcdev::", "")
    hdr = rows[a + 1]
    col = {h: i for i, h in enumerate(hdr)}
    body = [r for r in rows[a + 2:b] if len(r) == len(hdr)]
    n = lambda r: int(r[col["Instructions Executed"]])
    tot = sum(n(r) for r in body)
    if (name, tot) in seen or not tot:
        continue
    seen.add((name, tot))
    exits = [k for k, r in enumerate(body) if "EXIT" in r[col["Source"]]]
    sub = sum(n(r) for r in body[exits[-1] + 1:])
    calls = [(k, n(r), r[col["Avg. Threads Executed"]]) for k, r in enumerate(body) if "CALL" in r[col["Source"]] and n(r) > 0]
    print(f"{name:34s} warp-inst {tot:>11d}  in subroutines {100 * sub / tot:5.1f}%  call sites (sass index, warps, threads): {calls}")

# second pass: which of the out-of-line segments are the compiler's f64 division / sqrt special-case routines (short, no calls of
# their own, recognisable by the denormal-threshold compare / MUFU.RSQ64H) as opposed to this code's own __noinline__ functions
print()
seen = set()
for a, b in zip(secs[:-1], secs[1:]):
    name = rows[a][1].split("(rtcdev")[0].replace("void rtcdev::", "")
    hdr = rows[a + 1]
    col = {h: i for i, h in enumerate(hdr)}
    body = [r for r in rows[a + 2:b] if len(r) == len(hdr)]
    n = lambda r: int(r[col["Instructions Executed"]])
    tot = sum(n(r) for r in body)
    if (name, tot) in seen or not tot:
        continue
    seen.add((name, tot))
    exits = [k for k, r in enumerate(body) if "EXIT" in r[col["Source"]]]
    first, out = exits[-1] + 1, []
    for k in range(exits[-1] + 1, len(body)):
        if "RET" in body[k][col["Source"]]:
            seg = body[first:k + 1]
            text = " ".join(r[col["Source"]] for r in seg)
            kind = "div" if "1.469367938527859385e-39" in text and len(seg) < 130 else "sqrt" if "RSQ64H" in text and len(seg) < 80 else None
            if kind and sum(n(r) for r in seg):
                out.append((kind, round(100 * sum(n(r) for r in seg) / tot, 1), max(n(r) for r in seg)))
            first = k + 1
    print(f"{name:34s} warp-inst {tot:>11d}  special-case routines (kind, % of kernel, entries): {out}")
