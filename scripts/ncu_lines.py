"""Join an ncu SASS source page with nvdisasm line info and aggregate per CUDA source line.

    ncu -i rep.ncu-rep --page source --csv --kernel-id ::regex:<kernel>:<n> > page.csv
    cuobjdump -xelf all librt_b200.so ; nvdisasm -g -c kernels.sm_100a.cubin > kernels.sass
    python scripts/ncu_lines.py page.csv kernels.sass <mangled-name-substring> [top]
"""
import collections
import csv
import re
import sys

page, sass, func = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40

# 1. line info per instruction of the function, in order
lines, cur, inside = [], None, False
for ln in open(sass, errors="ignore"):
    if ln.lstrip().startswith(".section"):
        inside = (".text." in ln) and (func in ln)
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+\S", ln):
        lines.append(cur)

rows = list(csv.reader(open(page)))
hdr = rows[1]
body = [r for r in rows[2:] if len(r) == len(hdr)]
col = {h: i for i, h in enumerate(hdr)}
if len(body) != len(lines):
    print(f"warning: {len(body)} SASS rows in the page vs {len(lines)} in the disassembly", file=sys.stderr)
agg = collections.defaultdict(lambda: [0, 0, 0, 0])
tot = [0, 0, 0, 0]
for r, loc in zip(body, lines):
    v = [int(float(r[col["Instructions Executed"]] or 0)), int(float(r[col["Thread Instructions Executed"]] or 0)),
         int(float(r[col["Warp Stall Sampling (All Samples)"]] or 0)), 1]
    for k in range(4):
        agg[loc][k] += v[k]
        tot[k] += v[k]
print(f"total: warp-instr {tot[0]}  thread-instr {tot[1]}  avg active {tot[1] / max(tot[0], 1):.2f}  samples {tot[2]}  SASS {tot[3]}")
print(f"{'file:line':28s} {'warp-instr':>12s} {'share':>6s} {'active':>6s} {'samples':>8s} {'share':>6s} {'sass':>5s}")
for loc, v in sorted(agg.items(), key=lambda kv: -kv[1][2])[:top]:
    name = f"{loc[0]}:{loc[1]}" if loc else "?"
    print(f"{name:28s} {v[0]:12d} {v[0] / tot[0]:6.3f} {v[1] / max(v[0], 1):6.1f} {v[2]:8d} {v[2] / max(tot[2], 1):6.3f} {v[3]:5d}")
