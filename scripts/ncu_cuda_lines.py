"""Instructions executed and stall samples per CUDA source line: python scripts/ncu_cuda_lines.py <rep> <kernel-regex> [n]"""
import csv, subprocess, sys
rep, rx = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda", "--csv", "--kernel-name", "regex:" + rx, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
his = [i for i, r in enumerate(rows) if r and r[0] in ("Line #", "#")]
tot_ex = tot_s = 0
recs = []
fname = ""
for i, r in enumerate(rows):
    if r and r[0] == "File Path": fname = r[1] if len(r) > 1 else ""
    if r and r[0] in ("Line #", "#"):
        hdr = r
        isrc, isamp, iex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
        j = i + 1
        while j < len(rows) and rows[j] and rows[j][0] not in ("Line #", "#", "File Path", "Kernel Name"):
            rr = rows[j]
            try:
                ex, sm = int(rr[iex] or 0), int(rr[isamp] or 0)
            except Exception:
                ex = sm = 0
            if ex or sm: recs.append((fname.split("/")[-1], rr[0], ex, sm, rr[isrc].strip()[:100]))
            tot_ex += ex; tot_s += sm
            j += 1
print("total instructions", tot_ex, "samples", tot_s)
for f, ln, ex, sm, src in sorted(recs, key=lambda x: -x[2])[:n]:
    print("%-14s %5s ex=%9d %5.1f%%  samp=%5d %5.1f%%  %s" % (f, ln, ex, 100.0 * ex / max(1, tot_ex), sm, 100.0 * sm / max(1, tot_s), src))
