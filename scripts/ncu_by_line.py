"""Instructions executed / stall samples per CUDA source line of one kernel, by joining the ncu SASS page (execution order =
cubin order) with nvdisasm's line markers: python scripts/ncu_by_line.py <rep> <kernel-regex> <object.o> <mangled-substring> [n]"""
import collections, csv, re, subprocess, sys, tempfile, os, glob
rep, rx, obj, mangled = sys.argv[1:5]
n = int(sys.argv[5]) if len(sys.argv) > 5 else 40
d = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, capture_output=True)
cubin = glob.glob(d + "/*.cubin")[0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
# locate the function
lines = []   # (file:line) per instruction in order
cur = None; infn = False
for l in dis:
    m = re.match(r"\s*\.text\.(\S+):", l)
    if m: infn = mangled in m.group(1); continue
    if not infn: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", l): lines.append(cur)
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + rx, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]
isamp, iex = hdr.index("# Samples"), hdr.index("Instructions Executed")
seen = set(); sass = []
for r in rows[hi + 1:]:
    if len(r) <= iex or r[0] in seen or r[0] == "Address": continue
    seen.add(r[0])
    def num(x):
        try: return int(x)
        except Exception: return 0
    sass.append((num(r[iex]), num(r[isamp])))
print("sass instructions: ncu", len(sass), "nvdisasm", len(lines))
agg = collections.defaultdict(lambda: [0, 0, 0])
for (ex, sm), ln in zip(sass, lines):
    a = agg[ln]; a[0] += ex; a[1] += sm; a[2] += 1
tex = sum(a[0] for a in agg.values()); ts = sum(a[1] for a in agg.values())
src = {}
for (f, ln) in agg:
    if f and f not in src:
        p = [x for x in glob.glob("/root/repo/wormhole_b200/csrc/" + f)]
        src[f] = open(p[0]).read().splitlines() if p else []
print("total executed", tex, "samples", ts)
for (key, a) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:n]:
    f, ln = key if key else ("?", 0)
    text = src.get(f, [])[ln - 1].strip()[:90] if f in src and 0 < ln <= len(src[f]) else ""
    print("%-12s %4d ex=%9d %5.1f%% samp=%5d %5.1f%% sass=%4d | %s" % (f, ln, a[0], 100.0 * a[0] / max(1, tex), a[1], 100.0 * a[1] / max(1, ts), a[2], text))
