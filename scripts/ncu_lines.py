"""Source-line hot spots of one kernel from an .ncu-rep (needs -lineinfo + --import-source on).
usage: python scripts/ncu_lines.py <file.ncu-rep> [top N]"""
import csv, subprocess, sys, io
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], stdout=subprocess.PIPE,
                     stderr=subprocess.DEVNULL, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
cur, hdr, data = None, None, []
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or len(r) < len(hdr) or r[2] != "-":   # only the per-line aggregate rows (Address == "-")
        continue
    try:
        ns = int(r[hdr.index("# Samples")]); ni = int(r[hdr.index("Instructions Executed")])
        nt = int(r[hdr.index("Thread Instructions Executed")])
    except ValueError:
        continue
    data.append((ns, ni, nt, cur, r[0], r[1].strip()[:100]))
tot = sum(d[0] for d in data) or 1
toti = sum(d[1] for d in data) or 1
print(f"total samples {tot}, warp instructions {toti}")
for d in sorted(data, reverse=True)[:top]:
    print(f"{100 * d[0] / tot:5.1f}% smp {100 * d[1] / toti:5.1f}% inst  thr/inst {d[2] / max(d[1], 1):5.1f}  {d[3]}:{d[4]}  {d[5]}")
