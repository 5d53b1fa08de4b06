"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: total time, launches and share per kernel."""
import csv, sys, re, collections
rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith("==")]
rd = csv.DictReader(lines)
tot = collections.defaultdict(float); cnt = collections.Counter()
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum": continue
    v = float(r["Metric Value"].replace(",", "")); u = r["Metric Unit"]
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(u, 1.0)
    name = re.sub(r"\(T1.*|\((?:const |unsigned |float|int|TeleaWs|Cand|TempScal|Select).*", "", r["Kernel Name"])
    if not (len(sys.argv) > 2 and sys.argv[2] == "full"): name = re.sub(r"<.*", "", name)     # "full": keep template arguments
    tot[name] += v; cnt[name] += 1
T = sum(tot.values())
print(f"total {T/1e3:.2f} ms over {sum(cnt.values())} launches")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    print(f"{k[:58]:58s} {v/1e3:10.3f} ms {100*v/T:6.2f}%  n={cnt[k]:5d}  avg {v/cnt[k]:9.1f} us")
