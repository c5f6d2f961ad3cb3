"""Top SASS instructions by warp-stall samples for one kernel of an ncu report (read offline, no GPU).
usage: ncu_hotspots.py report.ncu-rep kernel_regex [launch_index_among_matches] [top_n]"""
import csv, io, subprocess, sys

rep, rx = sys.argv[1], sys.argv[2]
idx = int(sys.argv[3]) if len(sys.argv) > 3 else 0
top = int(sys.argv[4]) if len(sys.argv) > 4 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "-k", f"regex:{rx}"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
bounds = starts + [len(rows)]
sections = [rows[bounds[i]:bounds[i + 1]] for i in range(len(starts))]
if len(sections) >= 2 and len(sections) % 2 == 0 and sections[0] == sections[1]:
    sections = sections[::2]  # the source page lists every launch twice
print(f"{len(sections)} matching launches; showing #{idx}")
rows = sections[idx]
print(rows[0][1][:150])
hdr = rows[1]
col = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") and not h.endswith("_not_issued")]
body = [r for r in rows[2:] if len(r) == len(hdr)]
total = sum(int(r[col["# Samples"]] or 0) for r in body)
print(f"{len(body)} instructions, {total} samples")
agg = {}
for r in body:
    for h in stall_cols:
        v = int(r[col[h]] or 0)
        if v: agg[h] = agg.get(h, 0) + v
print("stall reasons:", ", ".join(f"{k[6:]} {100*v/max(total,1):.0f}%" for k, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
ranked = sorted(enumerate(body), key=lambda x: -int(x[1][col["# Samples"]] or 0))[:top]
for i, r in sorted(ranked):
    n = int(r[col["# Samples"]] or 0)
    why = sorted(((int(r[col[h]] or 0), h[6:]) for h in stall_cols), reverse=True)[:2]
    print(f"{i:5d} {100*n/max(total,1):5.1f}%  {r[col['Source']].strip()[:90]:90s} {' '.join(f'{w}:{c}' for c, w in why if c)}")
