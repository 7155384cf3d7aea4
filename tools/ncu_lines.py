"""Aggregate an `ncu --page source --csv --print-source cuda,sass` export per source line.
usage: python tools/ncu_lines.py export.csv '<substring of kernel name>' [top_n]"""
import csv, sys

rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
fn, fpath, agg = None, None, []
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fpath = r[1].split("/")[-1]
    elif r[0] == "Function Name":
        fn = r[1]
    elif r[0].isdigit() and fn and want in fn:
        try:
            agg.append((int(r[7]), int(r[6]), fpath, int(r[0]), r[1].strip()[:100]))
        except ValueError:
            pass
tot = sum(a[0] for a in agg)
smp = sum(a[1] for a in agg)
print("total warp-instructions", tot, "samples", smp)
for a in sorted(agg, reverse=True)[:top]:
    print(f"{a[0]:>10} {100 * a[0] / tot:5.1f}%  smp {100 * a[1] / max(smp, 1):5.1f}%  {a[2]}:{a[3]}  {a[4]}")
