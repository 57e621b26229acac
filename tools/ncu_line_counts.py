"""Per source line: executed warp instructions and stall samples from `ncu --page source --csv --print-source sass,cuda`.
usage: python tools/ncu_line_counts.py x.csv [units] [top]   (units = divisor for the instruction counts, e.g. warps x chunks)"""
import csv, collections, sys
def I(x):
    try: return int(x)
    except ValueError: return 0
rows = list(csv.reader(open(sys.argv[1])))
n = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
top = int(sys.argv[3]) if len(sys.argv) > 3 else 24
i = 0; tot = 0; per = {}; st = collections.Counter(); samp = {}; why = {}
while i < len(rows):
    if rows[i] and rows[i][0] == "Function Name":
        hdr = rows[i + 1]; c = hdr.index('Instructions Executed'); sc = hdr.index('# Samples'); f = rows[i - 1][1]
        col = {k: j for j, k in enumerate(hdr)}
        stall = [k for k in hdr if k.startswith("stall_") and "(" not in k]
        j = i + 2
        while j < len(rows) and rows[j] and rows[j][0] not in ("File Path", "Function Name"):
            r = rows[j]
            if r[0] != "" and f.endswith(('.cu', '.cuh', '.hpp', '.h')):
                v = I(r[c]); tot += v; k = (f[-14:], r[0], r[1].strip()[:64]); per[k] = v; samp[k] = I(r[sc])
                why[k] = sorted(((I(r[col[kk]]), kk[6:]) for kk in stall), reverse=True)[:2]
                for kk in stall: st[kk] += I(r[col[kk]])
            j += 1
        i = j
    else:
        i += 1
print("warp instructions", tot, "per unit %.1f" % (tot / n), "samples", sum(samp.values()))
s = sum(st.values())
print({k[6:]: "%.1f%%" % (100 * v / s) for k, v in st.most_common(8)})
print("-- by samples")
for k, v in sorted(samp.items(), key=lambda kv: -kv[1])[:top]: print("%6d smp %6.1f ins" % (v, per[k] / n), k, why[k])
