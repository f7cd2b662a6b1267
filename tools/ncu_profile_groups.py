#!/usr/bin/env python
"""Per-instruction execution profile of one kernel of an .ncu-rep, grouped into runs of equal execution count
(= loop bodies): where the instructions, samples and shared wavefronts go.
usage: tools/ncu_profile_groups.py rep kernel-substring [min-share-%]"""
import csv, io, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
minshare = float(sys.argv[3]) if len(sys.argv) > 3 else 0.4
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(io.StringIO(raw)))
names = [r[rr[0].index("Kernel Name")] for r in rr[2:]]
rows = list(csv.reader(io.StringIO(src)))
heads = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
ki = next(i for i, n in enumerate(names) if pat in n)
hdr = rows[heads[ki]]
end = heads[ki + 1] if ki + 1 < len(heads) else len(rows)
data = [r for r in rows[heads[ki] + 1:end] if len(r) == len(hdr)]
ix = {h: i for i, h in enumerate(hdr)}
def f(r, k):
    try: return float(r[ix[k]])
    except Exception: return 0.0
recs = [(n, f(r, "Instructions Executed") / 1e6, f(r, "# Samples"), f(r, "L1 Wavefronts Shared") / 1e6, r[ix["Source"]]) for n, r in enumerate(data)]
T = sum(r[1] for r in recs); S = sum(r[2] for r in recs); Wv = sum(r[3] for r in recs)
print(f"{names[ki][:70]}: {T:.0f}M warp instructions, {Wv:.0f}M shared wavefronts, {len(recs)} SASS lines")
groups = []
for r in recs:
    if groups and abs(groups[-1][2] - r[1]) <= 0.005 * max(r[1], 1e-9):
        g = groups[-1]; g[1] = r[0]; g[3] += 1; g[4] += r[2]; g[5] += r[3]; g[6] += ("LDS" in r[4])
    else:
        groups.append([r[0], r[0], r[1], 1, r[2], r[3], int("LDS" in r[4])])
for g in groups:
    w = g[2] * g[3]
    if 100 * w / T >= minshare:
        print(f"  sass {g[0]:5d}-{g[1]:5d} n={g[3]:4d} exec={g[2]:8.2f}M inst={w:8.1f}M ({100*w/T:4.1f}%) samples {100*g[4]/S:4.1f}% waves {g[5]:7.1f}M lds/iter={g[6]}")
print("  sections between barriers:")
cur = [0, 0.0, 0.0, 0.0]
for r in recs:
    cur[1] += r[1]; cur[2] += r[2]; cur[3] += r[3]
    if "BAR" in r[4] or r[0] == recs[-1][0]:
        print(f"    sass {cur[0]:5d}-{r[0]:5d}: inst {cur[1]:8.1f}M ({100*cur[1]/T:4.1f}%) samples {100*cur[2]/S:4.1f}% waves {cur[3]:7.1f}M")
        cur = [r[0] + 1, 0.0, 0.0, 0.0]
if "--dump" in sys.argv:
    for r in recs:
        print(f"{r[0]:5d} {r[1]:9.2f} {r[2]:7.0f} {r[3]:8.2f}  {r[4][:100]}")
