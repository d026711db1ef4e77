#!/usr/bin/env python
"""Aggregate an ncu report's per-SASS-instruction metrics by CUDA source line (needs -lineinfo).

    python tools/ncu_by_line.py REPORT.ncu-rep LIB.so MANGLED_SUBSTR [--top N]

ncu's CSV source page only lists SASS; nvdisasm --print-line-info gives the line of every SASS instruction of the
same cubin in the same order, so the two are joined by instruction index."""
import csv
import re
import subprocess
import sys
import tempfile
import os
from collections import defaultdict


def sass_lines(lib, mangled_sub):
    d = tempfile.mkdtemp()
    subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=d, stdout=subprocess.DEVNULL)
    for f in sorted(os.listdir(d)):
        txt = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(d, f)], capture_output=True, text=True).stdout
        if mangled_sub not in txt:
            continue
        out, on, cur = [], False, ("?", 0)
        for ln in txt.splitlines():
            if ln.startswith(".text."):
                on = mangled_sub in ln
                continue
            if not on:
                continue
            m = re.match(r'\s*//## File "(.*)", line (\d+)', ln)
            if m:
                cur = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
                out.append((cur, ln.split("*/", 1)[1].strip()))
            if ln.startswith("\t.section") and out:
                break
        if out:
            return out
    raise SystemExit("function not found")


def main():
    rep, lib, sub = sys.argv[1:4]
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 40
    csvtxt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(csvtxt.splitlines()))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    col = {n: hdr.index(n) for n in ("Source", "# Samples", "Instructions Executed", "Thread Instructions Executed")}
    stall_cols = [(n, i) for i, n in enumerate(hdr) if n.startswith("stall_") and "Not Issued" not in n]
    body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
    sl = sass_lines(lib, sub)
    if len(sl) != len(body):
        print("warning: %d SASS rows in report vs %d in cubin (is the .so the one profiled?)" % (len(body), len(sl)), file=sys.stderr)
    agg = defaultdict(lambda: [0, 0, 0, defaultdict(int)])
    tot = [0, 0, 0]
    for r, (loc, _) in zip(body, sl):
        a = agg[loc]
        v = [int(float(r[col[k]] or 0)) for k in ("# Samples", "Instructions Executed", "Thread Instructions Executed")]
        for i in range(3):
            a[i] += v[i]; tot[i] += v[i]
        for n, i in stall_cols:
            if r[i] and r[i] != "0":
                a[3][n] += int(float(r[i]))
    print("total samples %d, warp-inst %d, thread-inst %d, avg threads/inst %.2f" % (tot[0], tot[1], tot[2], tot[2] / max(tot[1], 1)))
    byfile = defaultdict(lambda: [0, 0, 0])
    for (f, l), a in agg.items():
        for i in range(3):
            byfile[f][i] += a[i]
    for f, a in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
        print("  %-28s samples %5.1f%%  warp-inst %5.1f%%  thr/inst %5.2f" % (f, 100 * a[0] / tot[0], 100 * a[1] / tot[1], a[2] / max(a[1], 1)))
    print("%-28s %7s %7s %8s  top stalls" % ("file:line", "samp%", "inst%", "thr/inst"))
    for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        st = ", ".join("%s %d" % (n[6:], c) for n, c in sorted(a[3].items(), key=lambda kv: -kv[1])[:3])
        print("%-28s %6.2f%% %6.2f%% %8.2f  %s" % ("%s:%d" % (f, l), 100 * a[0] / tot[0], 100 * a[1] / tot[1], a[2] / max(a[1], 1), st))


if __name__ == "__main__":
    main()
