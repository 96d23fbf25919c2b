"""summarise `ncu --page source --csv` output: per kernel, the source lines / SASS with the most stall samples.
usage: ncu -i X.ncu-rep --page source --csv > src.csv; python scripts/ncu_hot.py src.csv [kernel-substring] [top]"""
import csv, sys

def main():
    path = sys.argv[1]; want = sys.argv[2] if len(sys.argv) > 2 else ""; top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    rows = list(csv.reader(open(path)))
    starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"] + [len(rows)]
    for a, b in zip(starts[:-1], starts[1:]):
        name = rows[a][1]
        if want not in name:
            continue
        h = rows[a + 1]; data = [r for r in rows[a + 2:b] if len(r) == len(h)]
        if not data:
            continue
        iS = h.index("Source"); iN = h.index("# Samples"); iE = h.index("Instructions Executed")
        is_sass = "Address" in h
        tot = sum(int(r[iN]) for r in data if r[iN].isdigit()) or 1
        tote = sum(int(r[iE]) for r in data if r[iE].isdigit())
        print(f"=== {name[:60]} [{'SASS' if is_sass else 'source'}] samples={tot} inst={tote}")
        order = sorted(range(len(data)), key=lambda i: -(int(data[i][iN]) if data[i][iN].isdigit() else 0))[:top]
        for i in sorted(order):
            r = data[i]
            print(f"{i:5d} {100.0 * int(r[iN]) / tot:5.1f}% inst={r[iE]:>10s}  {r[iS].strip()[:120]}")

main()
