import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
hdr, data = rows[hi], rows[hi + 1:]
ki, mi, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.defaultdict(lambda: collections.defaultdict(float))
cnt = collections.Counter()
for r in data:
    if len(r) <= vi:
        continue
    name = r[ki].split("(")[0][:56]
    v = float(r[vi].replace(",", ""))
    u = r[ui]
    if r[mi].startswith("gpu__time"):
        v = v / 1000 if u == "ns" else (v * 1000 if u == "ms" else v)
        cnt[name] += 1
    else:
        v = v * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1, "Gbyte": 1e3}.get(u, 1)
    agg[name][r[mi]] += v
tot = sum(a["gpu__time_duration.sum"] for a in agg.values())
print("%-58s %5s %9s %7s %9s %9s" % ("kernel", "n", "avg us", "share", "rd MB/l", "wr MB/l"))
for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["gpu__time_duration.sum"])[:28]:
    n = cnt[k]
    print("%-58s %5d %9.2f %6.1f%% %9.2f %9.2f" % (k, n, a["gpu__time_duration.sum"] / n, 100 * a["gpu__time_duration.sum"] / tot,
                                                a["dram__bytes_read.sum"] / n, a["dram__bytes_write.sum"] / n))
