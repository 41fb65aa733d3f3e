"""Summarise an ncu report's SASS page for one kernel: instructions per tile by execution-count class and opcode,
top stall sites.  usage: python tools/sass_profile.py report.ncu-rep tiles [dump_from dump_to]"""
import collections
import csv
import io
import re
import subprocess
import sys

rep, tiles = sys.argv[1], float(sys.argv[2])
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"rows": []}
        blocks.append(cur)
    elif r and r[0] == "Address":
        cur["hdr"] = r
    elif cur is not None and r:
        cur["rows"].append(r)
b = blocks[0]
h = b["hdr"]
ia, ie, iss = h.index("Source"), h.index("Instructions Executed"), h.index("Warp Stall Sampling (All Samples)")
R = b["rows"]
c, n, ops = collections.Counter(), collections.Counter(), collections.Counter()
for r in R:
    e = int(r[ie])
    c[e] += e
    n[e] += 1
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_]+(\.[A-Z0-9_.]+)?)", r[ia])
    ops[m.group(2) if m else r[ia][:8]] += e
tot = sum(c.values())
print("static instructions %d, executed per tile %.1f" % (len(R), tot / tiles))
for e, v in sorted(c.items(), key=lambda kv: -kv[1])[:8]:
    print("  executed %.2f x per tile: %4d static, %5.1f%%, %7.1f per tile" % (e / tiles, n[e], 100 * v / tot, v / tiles))
for op, v in ops.most_common(24):
    print("  %-22s %5.1f%% %7.1f/tile" % (op, 100 * v / tot, v / tiles))
stt = sum(int(r[iss]) for r in R)
for i in sorted(range(len(R)), key=lambda i: -int(R[i][iss]))[:8]:
    print("  stall %5.2f%%  [%d] %s" % (100 * int(R[i][iss]) / stt, i, R[i][ia].strip()[:70]))
if len(sys.argv) > 4:
    for i in range(int(sys.argv[3]), int(sys.argv[4])):
        print(i, R[i][ie].rjust(9), R[i][ia].strip()[:100])
