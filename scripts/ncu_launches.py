"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel name.
  python scripts/ncu_launches.py launches.csv ["header line"] [--step SUBSTR]
--step: keep exactly one step = the launches after the second-to-last launch whose name contains SUBSTR up to and including
the last one (the optimizer kernel closes a training step)."""
import csv, re, sys
step = None
if "--step" in sys.argv:
    i = sys.argv.index("--step")
    step = sys.argv[i + 1]
    del sys.argv[i:i + 2]
rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
body = rows[1:]
if step:
    marks = [i for i, r in enumerate(body) if step in r[ki]]
    if len(marks) >= 2:
        body = body[marks[-2] + 1:marks[-1] + 1]
agg = {}
for r in body:
    v = float(r[vi].replace(",", ""))
    v = v / 1e3 if r[ui] in ("ns", "nsecond") else (v * 1e3 if r[ui] in ("ms", "msecond") else v)
    name = re.sub(r"^void |\(.*$|hn::|\(anonymous namespace\)::", "", r[ki])[:80]
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
if len(sys.argv) > 2:
    print(sys.argv[2])
print(f"{sum(a[0] for a in agg.values())} launches, {tot:.1f} us total (cold-cache, serialised: compare shares)")
for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:80s} n={n:4d} {us:9.1f} us {100 * us / tot:5.1f}%  avg {us / n:7.1f} us")
