"""Hot SASS lines of the first kernel in an ncu source page:  python scripts/ncu_hot_sass.py page.csv [min_share]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
share = float(sys.argv[2]) if len(sys.argv) > 2 else 0.004
hdr = rows[1]
col = {h: i for i, h in enumerate(hdr)}
body = []
for r in rows[2:]:
    if len(r) != len(hdr) or r[col["Instructions Executed"]] == "Instructions Executed":
        if body:
            break
        continue
    body.append(r)
tot = sum(int(float(r[col["Instructions Executed"]] or 0)) for r in body)
samp = sum(int(float(r[col["Warp Stall Sampling (All Samples)"]] or 0)) for r in body)
print(f"{len(body)} SASS, {tot} warp-instr, {samp} samples")
for k, r in enumerate(body):
    n = int(float(r[col["Instructions Executed"]] or 0))
    s = int(float(r[col["Warp Stall Sampling (All Samples)"]] or 0))
    if n > tot * share or s > samp * share * 2:
        print(f'{k:5d} {n:9d} {s:5d} {r[col["Avg. Threads Executed"]]:>5s}  {r[col["Source"]][:100]}')
