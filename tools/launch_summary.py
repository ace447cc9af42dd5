"""Per-kernel totals of an ncu launch list (`--metrics gpu__time_duration.sum --csv`):
    python tools/launch_summary.py gpurun_out/launches.csv > profiles/<round>_launch_summary.txt"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
h = rows[hi]
kn, mv = h.index("Kernel Name"), h.index("Metric Value")
tot, cnt, big = collections.defaultdict(float), collections.Counter(), []
for r in rows[hi + 1:]:
    if len(r) <= mv:
        continue
    try:
        v = float(r[mv].replace(",", ""))
    except ValueError:
        continue
    tot[r[kn]] += v
    cnt[r[kn]] += 1
    if "step_many" in r[kn] and v > 80000:
        big.append(v)
s = sum(tot.values())
print("ncu --metrics gpu__time_duration.sum --clock-control none  python bench.py --steps 20 --warmup 3 --no-cpu-baseline")
print("Per-launch times under ncu are cold-cache and serialised: shares, not absolutes.  The timed region of the headline is 20")
print(f"launches of step_many_kernel<0,0> with T = 32 and nothing else (the {len(big)} launches above 80 us: warm-up + timed + e2e legs;")
print(f"mean {sum(big) / max(len(big), 1) / 1e3:.1f} us under ncu, to compare with roofline.avg_launch_ms of the bench line), so the kernel's share of a")
print("bench step is 100 %.  The other launches belong to the extras (T = 1 steps, frames) and to the DP section (enumeration")
print("levels, sweeps, cooperative solves, the batched-layouts leg; the list is cut at 4,000 launches).\n")
for k, v in sorted(tot.items(), key=lambda x: -x[1])[:18]:
    print(f"{v / 1e3:10.1f} us {100 * v / s:5.1f}% n={cnt[k]:5d} avg {v / cnt[k] / 1e3:8.2f} us  {k[:100]}")
