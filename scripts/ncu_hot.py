"""Hot spots of an .ncu-rep captured with --import-source on: opcode mix (instructions vs stall samples) and the
SASS lines with the most samples plus their two main stall reasons.
  python scripts/ncu_hot.py file.ncu-rep [N]"""
import csv, io, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = rows[1]
si, ei, st = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
cols = {k: h.index(k) for k in h if k.startswith("stall_") and "Not Issued" not in k}
tot = ti = 0
L, agg = [], {}
for r in rows[2:]:
    try:
        s, n = float(r[st].replace(",", "") or 0), float(r[ei].replace(",", "") or 0)
    except ValueError:
        continue
    tot += s; ti += n; L.append((s, r))
    t = r[si].split()
    op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
    a = agg.setdefault(op, [0, 0]); a[0] += n; a[1] += s
print(rows[0][1][:110])
print(f"samples {tot:.0f}  warp instructions {ti:.3e}")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:12]:
    print(f"  {k:10s} inst {100 * v[0] / ti:5.1f}%  samples {100 * v[1] / tot:5.1f}%")
reasons = {k: sum(float(r[i].replace(",", "") or 0) for _, r in L) for k, i in cols.items()}
print("  stall reasons:", ", ".join(f"{k[6:]} {100 * v / tot:.0f}%" for k, v in sorted(reasons.items(), key=lambda kv: -kv[1])[:6]))
L.sort(key=lambda x: -x[0])
for s, r in L[:int(sys.argv[2]) if len(sys.argv) > 2 else 12]:
    top = sorted(((float(r[i].replace(",", "") or 0), k[6:]) for k, i in cols.items()), reverse=True)[:2]
    print(f"  {100 * s / tot:5.1f}%  {r[si][:64]:64s} {top}")
