"""Print the CTA-0 timeline dumped by PIPER_B200_MRF_TRACE=<path> (debug aid for conv_mrf.cu).
usage: python tools/mrf_trace.py <path>.<C> [n_lo n_hi]"""
import sys
from collections import defaultdict
TAG = {1: "c1.beg", 2: "c1.go", 3: "c1.end", 5: "c2.beg", 6: "c2.go", 7: "c2.end", 10: "ep.beg", 11: "ep.s_ok", 12: "ep.d_ok",
       13: "ep.end", 14: "out.beg", 15: "out.go", 16: "out.end", 20: "x.wait", 21: "x.load"}
ev = []
for ln in open(sys.argv[1]):
    C, m, w, tag, n, i, c = map(int, ln.split())
    ev.append((w, tag, n, i, c))
n_lo = int(sys.argv[2]) if len(sys.argv) > 2 else 6
n_hi = int(sys.argv[3]) if len(sys.argv) > 3 else 8
# unwrap the 32-bit clock per warp
by_w = defaultdict(list)
for e in ev: by_w[e[0]].append(e)
out = []
for w, lst in by_w.items():
    base = 0; prev = None
    for (w_, tag, n, i, c) in lst:
        if prev is not None and c < prev: base += 1 << 32
        prev = c
        out.append((c + base, w, tag, n, i))
out.sort()
t0 = None
sel = [e for e in out if n_lo <= e[3] <= n_hi and e[2] < 20 and (e[1] in (1, 2, 3, 7))]
if sel: t0 = sel[0][0]
for (c, w, tag, n, i) in sel:
    col = {1: 0, 2: 1, 3: 2, 7: 3}[w]
    print("%8d  %s%-8s n=%d i=%d" % (c - t0, " " * (24 * col), TAG.get(tag, str(tag)), n, i))
# summary: per-phase durations
def dur(w, a, b):
    st = {}; tot = 0; cnt = 0
    for (c, w_, tag, n, i) in out:
        if w_ != w: continue
        if tag == a: st[(n, i)] = c
        elif tag == b and (n, i) in st: tot += c - st.pop((n, i)); cnt += 1
    return tot / max(cnt, 1), cnt
span = out[-1][0] - out[0][0]
print("total span cycles", span)
# warps: 0 producer, 1/2 conv1 issuers (even/odd tiles), 3/4 conv2 issuers, 5.. epilogue (group g = warps 5+4g .. 8+4g)
warps = sorted(by_w)
PAIRS = [("c1 wait", 1, 2), ("c1 issue", 2, 3), ("c2 wait", 5, 6), ("c2 issue", 6, 7), ("w_full wait", 24, 25),
         ("x wait", 20, 21), ("w_empty wait", 22, 23), ("ep s wait", 10, 11), ("ep d wait", 11, 12), ("ep work", 12, 13),
         ("out wait", 14, 15), ("out work", 15, 16)]
for w in warps:
    if w > 9 and w != warps[-1]:
        continue
    for name, a_, b_ in PAIRS:
        d, c = dur(w, a_, b_)
        if c:
            print("warp %2d %-14s avg %8.0f cycles x %5d = %5.1f%% of span" % (w, name, d, c, 100.0 * d * c / span))
