"""Copy one profiling pass (scripts/gpu_profile.sh <tag>) from gpurun_out/ into profiles/ under a round name and
derive profiles/traffic.json (dram bytes per launch, read by bench.py's roofline.traffic).
usage: python scripts/make_profiles.py <tag> <round-name>"""
import csv, json, os, re, shutil, sys
tag, rnd = sys.argv[1], sys.argv[2]
G, P = "gpurun_out", "profiles"
os.makedirs(P, exist_ok=True)
shutil.copy(f"{G}/{tag}_launches.csv", f"{P}/{rnd}_launches.csv")
shutil.copy(f"{G}/{tag}_plain.json", f"{P}/{rnd}_bench_plain.json")
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
traffic = {}
for f in sorted(os.listdir(G)):
    m = re.match(rf"{tag}_(k_\w+)_raw\.csv$", f)
    if not m:
        continue
    name = m.group(1)
    rows = list(csv.reader(open(f"{G}/{f}")))
    hdr, units = rows[0], rows[1]
    ik, ir, iw, it = (hdr.index(x) for x in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum"))
    for r in rows[2:]:
        kn = r[ik]
        kn = re.sub(r"^void ", "", kn)
        short = kn.split("(")[0]
        b = float(r[ir]) * UNIT[units[ir]] + float(r[iw]) * UNIT[units[iw]]
        traffic.setdefault(short, []).append({"dram_bytes": b, "ncu_ms": float(r[it]) * {"us": 1e-3, "ms": 1, "ns": 1e-6}[units[it]]})
    with open(f"{P}/{rnd}_{name}.txt", "w") as o:
        o.write(f"# ncu --set full --clock-control none, kernel {name}, workload configs[1] (bench.py --steps 1 --warmup 3); one block per launch\n")
        o.write(open(f"{G}/{tag}_{name}_summary.txt").read())
        o.write("\n# source-line hot spots (samples, warp instructions, threads per instruction)\n")
        o.write(open(f"{G}/{tag}_{name}_lines.txt").read())
out = {}
for k, v in traffic.items():
    out[k] = {"dram_bytes_per_launch": sum(x["dram_bytes"] for x in v) / len(v), "launches_captured": len(v),
              "ncu_ms_per_launch": sum(x["ncu_ms"] for x in v) / len(v)}
json.dump({"source": f"{rnd} ncu --set full captures, configs[1]", "kernels": out}, open(f"{P}/traffic.json", "w"), indent=1)
print(json.dumps(out, indent=1))
