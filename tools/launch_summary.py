"""Folds an ncu launch list (--metrics gpu__time_duration.sum --csv) into time and share per kernel.
    python tools/launch_summary.py gpurun_out/r2_launches.csv "command that was profiled" > profiles/r2_launches_summary.txt"""
import csv
import sys
from collections import defaultdict

path = sys.argv[1]
what = sys.argv[2] if len(sys.argv) > 2 else "?"
rows = [r for r in csv.reader(open(path, errors="replace")) if len(r) > 14 and r[12] == "gpu__time_duration.sum"]
tot, cnt = defaultdict(float), defaultdict(int)
for r in rows:
    name = r[4].replace("ssb::", "").replace("(anonymous namespace)::", "").replace("<unnamed>::", "")
    scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[13], 1e-6)
    tot[name] += float(r[14].replace(",", "")) * scale
    cnt[name] += 1
total = sum(tot.values()) or 1.0
print(f"ncu launch list of `{what}`;")
print("per-launch times are cold-cache and serialised: the SHARES are what compares with bench.py's `kernels[].share_of_step`")
print(f"total {total:.2f} ms over {sum(cnt.values())} launches")
for name, t in sorted(tot.items(), key=lambda kv: -kv[1]):
    print(f"{t:9.3f} ms {cnt[name]:5d} launches {100 * t / total:5.1f}%  {name[:100]}")
