"""Aggregate an `ncu --page source --csv --print-source sass,cuda` export by CUDA source line.
usage: ncu -i rep --page source --csv --print-source sass,cuda > x.csv; python tools/ncu_source_lines.py x.csv [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
def num(x):
    try: return int(x)
    except ValueError: return 0
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
i = 0
while i < len(rows):
    if rows[i] and rows[i][0] == "Function Name":
        name = rows[i][1]; hdr = rows[i + 1]; col = {n: j for j, n in enumerate(hdr)}
        j = i + 2; lines = []
        while j < len(rows) and rows[j] and rows[j][0] not in ("File Path", "Function Name"):
            r = rows[j]
            if r[0] != "":
                lines.append(r)
            j += 1
        tot = sum(num(r[col["# Samples"]]) for r in lines)
        print("==", rows[i-1][1][-40:] if i else "", name[:90], "samples", tot)
        stall = [k for k in hdr if k.startswith("stall_") and "(" not in k]
        for r in sorted(lines, key=lambda r: -num(r[col["# Samples"]]))[:top]:
            st = sorted(((num(r[col[k]]), k) for k in stall), reverse=True)[:2]
            print("%5s %6.2f%% %-90s %s" % (r[0], 100.0 * num(r[col["# Samples"]]) / max(tot, 1), r[1].strip()[:90], st))
        i = j
    else:
        i += 1
