"""Summarises an ncu --csv launch list: the launches of the LAST step in order, with duration and
DRAM bytes when present.   python tools/launch_summary.py launches.csv [n_last]"""
import collections
import csv
import sys

path = sys.argv[1]
nlast = int(sys.argv[2]) if len(sys.argv) > 2 else 0
lines = [l for l in open(path) if not l.startswith("==")]
rows = list(csv.DictReader(lines))
by_id = collections.OrderedDict()
for r in rows:
    d = by_id.setdefault(r["ID"], {"name": r["Kernel Name"]})
    v = float(r["Metric Value"].replace(",", ""))
    u = r["Metric Unit"]
    m = r["Metric Name"]
    if m.startswith("gpu__time"):
        d["us"] = v / 1e3 if u in ("ns", "nsecond") else (v * 1e3 if u in ("ms", "msecond") else v)
    else:
        mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
        d[m] = v * mult
items = list(by_id.values())
if nlast:
    items = items[-nlast:]
tot = sum(i.get("us", 0) for i in items)
for i in items:
    rd, wr = i.get("dram__bytes_read.sum", 0), i.get("dram__bytes_write.sum", 0)
    gbs = (rd + wr) / (i["us"] * 1e-6) / 1e9 if i.get("us") else 0
    print(f"{i['us']:9.1f} us {100 * i['us'] / tot:5.1f}%  rd {rd / 1e6:8.2f} MB wr {wr / 1e6:8.2f} MB {gbs:7.0f} GB/s  {i['name'][:100]}")
print(f"{tot:9.1f} us total over {len(items)} launches")
