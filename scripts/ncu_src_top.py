"""Top SASS lines of one kernel of an ncu report by stall samples: python scripts/ncu_src_top.py <rep> <kernel-regex> [n]"""
import csv, subprocess, sys
rep, rx = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + rx, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]
ia, isamp, iex, il2 = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("L2 Theoretical Sectors Global")
def num(x):
    try: return int(x)
    except Exception: return 0
data = [r for r in rows[hi + 1:] if len(r) > il2 and r[0] != "Address"]
tot = sum(num(r[isamp]) for r in data)
print("total samples", tot, "instructions executed", sum(num(r[iex]) for r in data), "sass lines", len(data))
for r in sorted(data, key=lambda r: -num(r[isamp]))[:n]:
    print("%6s %5.1f%% ex=%8s l2=%8s  %s" % (r[isamp], 100.0 * num(r[isamp]) / max(1, tot), r[iex], r[il2], r[ia][:120]))
