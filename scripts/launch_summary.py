"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) of scripts/profile_step.py: the LAST complete
score step (launches from the last temb_mlp_kernel on), per kernel: launches, total us, share.
usage: python scripts/launch_summary.py gpurun_out/x_launches.csv > profiles/x_launch_summary.txt"""
import collections
import csv
import io
import sys

rows = [l for l in open(sys.argv[1]) if l.startswith('"')]
r = list(csv.reader(io.StringIO("".join(rows))))
h, d = r[0], r[1:]
ki, vi, gi = h.index("Kernel Name"), h.index("Metric Value"), h.index("Grid Size")
starts = [i for i, x in enumerate(d) if "temb_mlp_kernel" in x[ki]]
step = d[starts[-1]:]
agg = collections.defaultdict(lambda: [0, 0.0])
for x in step:
    k = x[ki].split("(")[0].replace("void ", "").replace("ditsep::", "")
    agg[k][0] += 1
    agg[k][1] += float(x[vi]) / 1e3
tot = sum(v[1] for v in agg.values())
print(f"# {sys.argv[1]}: {len(d)} launches captured, {len(starts)} score steps; last step: {len(step)} launches, "
      f"sum of per-launch durations (serialised, cold cache) {tot / 1e3:.3f} ms")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:64]:64s} n={v[0]:4d} {v[1]:9.1f} us {100 * v[1] / tot:5.1f} %")
