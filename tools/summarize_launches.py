"""Summarise an ncu --csv launch list (gpu__time_duration.sum) per kernel: count, average, share."""
import collections, csv, sys
path = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/launches.csv"
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
agg = collections.defaultdict(lambda: [0, 0.0])
for x in csv.DictReader(lines):
    try:
        v = float(x["Metric Value"].replace(",", ""))
    except (ValueError, KeyError):
        continue
    k = x["Kernel Name"][:56]
    agg[k][0] += 1
    agg[k][1] += v
tot = sum(v[1] for v in agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-58s n=%4d avg=%9.1f ns share=%.3f" % (k, v[0], v[1] / v[0], v[1] / tot))
