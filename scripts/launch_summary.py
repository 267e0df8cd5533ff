"""Summarise an ncu --csv launch list (gpu__time_duration.sum [+ lane/instruction metrics]) per kernel
and per wavefront iteration. usage: launch_summary.py launches.csv [every]"""
import csv, gzip, sys, collections

rows = list(csv.reader(gzip.open(sys.argv[1], "rt") if sys.argv[1].endswith(".gz") else open(sys.argv[1])))
every = int(sys.argv[2]) if len(sys.argv) > 2 else 10
start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
hdr = rows[start]
ik, im, iv, iid, ig = (hdr.index(k) for k in ("Kernel Name", "Metric Name", "Metric Value", "ID", "Grid Size"))
launches = collections.OrderedDict()
for r in rows[start + 1:]:
    if len(r) <= iv or not r[iid].isdigit():
        continue
    d = launches.setdefault(int(r[iid]), {"k": r[ik].split("(")[0].replace("void ", "").replace("vpt::", ""),
                                          "grid": int(r[ig].strip("()").split(",")[0])})
    d[r[im]] = float(r[iv].replace(",", ""))
tot = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0, 0.0])
for d in launches.values():
    t = tot[d["k"]]
    t[0] += 1
    t[1] += d.get("gpu__time_duration.sum", 0.0)
    inst = d.get("smsp__inst_executed.sum", 0.0)
    t[2] += inst
    t[3] += inst * d.get("smsp__thread_inst_executed_per_inst_executed.ratio", 0.0)
    t[4] += d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
allt = sum(t[1] for t in tot.values())
print(f"{len(launches)} launches, {allt / 1e6:.2f} ms of kernel time (serialised, cold cache, under ncu)")
print(f"{'kernel':28s} {'n':>5s} {'ms':>9s} {'share':>7s} {'us/launch':>10s} {'lanes/inst':>10s} {'dram/launch':>12s}")
for k, t in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:28s} {t[0]:5d} {t[1] / 1e6:9.2f} {100 * t[1] / allt:6.1f}% {t[1] / t[0] / 1e3:10.1f} "
          f"{(t[3] / t[2] if t[2] else 0):10.2f} {t[4] / t[0]:12.3g}")
print("\nper iteration (every %d): extend grid, us extend/shade/lightpdf, lanes extend/shade/lightpdf" % every)
it, cur = [], None
for d in launches.values():
    if "extend" in d["k"]:
        cur = {"g": d["grid"]}
        it.append(cur)
    if cur is None:
        continue
    key = "e" if "extend" in d["k"] else "s" if "shade" in d["k"] else "p" if "lightpdf" in d["k"] else None
    if key:
        cur[key] = d.get("gpu__time_duration.sum", 0.0) / 1e3
        cur[key + "l"] = d.get("smsp__thread_inst_executed_per_inst_executed.ratio", 0.0)
for i, c in enumerate(it):
    if i % every == 0:
        print(f"{i:5d} grid {c['g']:5d}  us {c.get('e', 0):8.1f} {c.get('s', 0):8.1f} {c.get('p', 0):8.1f}   "
              f"lanes {c.get('el', 0):5.2f} {c.get('sl', 0):5.2f} {c.get('pl', 0):5.2f}")
