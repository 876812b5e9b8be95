"""Per-launch DRAM traffic by kernel class from an .ncu-rep (`ncu --set full`), for bench.py's `roofline.traffic`.

  python tools/ncu_traffic.py extract REP.ncu-rep OUT.json "description of the capture"     (on the box or here; needs ncu, no GPU)
  python tools/ncu_traffic.py merge CONFIG_ID IN.json                                        (here: updates profiles/roofline_traffic.json)

The entry records the hash of the kernel sources it was captured on (bench.kernel_source_hash): bench.py reports the
traffic only if the kernels it times are the ones that were captured."""
import csv, io, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
CLASSES = {"trace_spawn": "trace_spawn", "trace_shadow": "trace_shadow", "light": "light"}
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def extract(rep, out, source):
    import bench

    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ik, ir, iw, it = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("gpu__time_duration.sum")
    per = {}
    for r in data:
        name = r[ik].split("(")[0].replace("void ", "").split("<")[0].split("::")[-1]
        cls = CLASSES.get(name)
        if cls is None:
            continue
        b = float(r[ir].replace(",", "")) * UNIT[units[ir]] + float(r[iw].replace(",", "")) * UNIT[units[iw]]
        e = per.setdefault(cls, {"bytes": 0.0, "launches": 0})
        e["bytes"] += b
        e["launches"] += 1
    entry = {"source": source, "kernel_source_hash": bench.kernel_source_hash(),
             "per_launch_dram_bytes": {k: v["bytes"] / v["launches"] for k, v in per.items()},
             "launches": {k: v["launches"] for k, v in per.items()}}
    with open(out, "w") as f:
        json.dump(entry, f, indent=1)
    print(json.dumps(entry))


def merge(config_id, path):
    tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    table = {}
    if os.path.exists(tpath):
        with open(tpath) as f:
            table = json.load(f)
        if "per_launch_dram_bytes" in table:  # the round-1 single-entry form
            table = {}
    with open(path) as f:
        table[config_id] = json.load(f)
    with open(tpath, "w") as f:
        json.dump(table, f, indent=1)
    print("profiles/roofline_traffic.json:", {k: v["kernel_source_hash"] for k, v in table.items()})


if __name__ == "__main__":
    if sys.argv[1] == "extract":
        extract(sys.argv[2], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else sys.argv[2])
    else:
        merge(sys.argv[2], sys.argv[3])
