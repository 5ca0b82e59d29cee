"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) per kernel: python scripts/launch_summary.py in.csv out.txt "command" """
import csv, sys, collections
src, out, cmd = sys.argv[1], sys.argv[2], (sys.argv[3] if len(sys.argv) > 3 else "")
rows = [r for r in csv.reader(open(src, errors="replace")) if len(r) > 5]
hdr = next(r for r in rows if "Kernel Name" in r)
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.OrderedDict()
unit = "ns"
for r in rows:
    if r is hdr or r[ki] == "Kernel Name":
        continue
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    unit = r[ui]
    scale = {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1.0)
    name = r[ki].split("(")[0][:52]
    c, t = agg.get(name, (0, 0.0))
    agg[name] = (c + 1, t + v * scale)
tot = sum(t for _, t in agg.values())
with open(out, "w") as f:
    f.write("ncu --metrics gpu__time_duration.sum --clock-control none: %s\n(cold-cache, serialised: compare SHARES)  unit=ns\n" % cmd)
    for name, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write("%-52s launches %4d  total %14.1f  share %6.2f%%  avg %12.1f\n" % (name, c, t, 100 * t / tot, t / c))
print(open(out).read())
