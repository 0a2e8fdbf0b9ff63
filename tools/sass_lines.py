"""Join an ncu SASS-page CSV (ncu -i X.ncu-rep --page source --csv --print-source sass) with the line table of the
same kernel from `nvdisasm -g -c -fun <symbol index> <cubin>`, by instruction order, and aggregate executed warp
instructions / stall samples per source line (innermost inlined location) and per file.

    python tools/sass_lines.py ncu_sass.csv[.gz] nvdisasm.txt [top]
"""
import csv
import gzip
import re
import sys
from collections import defaultdict


def read_ncu(path):
    op = gzip.open if path.endswith(".gz") else open
    rows = []
    with op(path, "rt") as fh:
        rd = csv.reader(fh)
        hdr = None
        for r in rd:
            if hdr is None:
                if r and r[0] == "Address":
                    hdr = r
                continue
            rows.append(r)
    ix = {n: i for i, n in enumerate(hdr)}
    return rows, ix


def read_disasm(path):
    """-> list of (file, line, text) per instruction in order"""
    out = []
    cur = ("?", 0)
    pat = re.compile(r'//## File "([^"]+)", line (\d+)')
    ins = re.compile(r"^\s+/\*([0-9a-f]{4,})\*/\s+(.*?);")
    for ln in open(path):
        m = pat.search(ln)
        if m:
            # the first File comment of a group is the innermost location; "inlined at" lines follow
            if "inlined at" not in ln:
                cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = ins.match(ln)
        if m:
            out.append((cur[0], cur[1], m.group(2)))
    return out


def main():
    rows, ix = read_ncu(sys.argv[1])
    dis = read_disasm(sys.argv[2])
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    print("ncu instructions %d, nvdisasm instructions %d" % (len(rows), len(dis)))
    n = min(len(rows), len(dis))
    by_line = defaultdict(lambda: [0, 0, 0, 0])
    by_file = defaultdict(lambda: [0, 0, 0, 0])
    tot = [0, 0, 0]
    for i in range(n):
        r = rows[i]
        ex = int(r[ix["Instructions Executed"]] or 0)
        th = int(r[ix["Thread Instructions Executed"]] or 0)
        sm = int(r[ix["# Samples"]] or 0)
        f, l, _ = dis[i]
        for d, k in ((by_line, (f, l)), (by_file, f)):
            d[k][0] += ex; d[k][1] += th; d[k][2] += sm; d[k][3] += 1
        tot[0] += ex; tot[1] += th; tot[2] += sm
    print("total warp instr %d, thread instr %d (%.1f lanes), samples %d" % (tot[0], tot[1], tot[1] / max(tot[0], 1), tot[2]))
    print("\nper file: file, warp-instr %, lanes, samples %, static SASS")
    for k, v in sorted(by_file.items(), key=lambda kv: -kv[1][0]):
        print("  %-18s %5.1f%%  %5.1f  %5.1f%%  %6d" % (k, 100.0 * v[0] / tot[0], v[1] / max(v[0], 1), 100.0 * v[2] / max(tot[2], 1), v[3]))
    print("\nper line: file:line, warp-instr %, lanes, samples %, static SASS")
    for k, v in sorted(by_line.items(), key=lambda kv: -kv[1][0])[:top]:
        print("  %-18s:%-5d %5.1f%%  %5.1f  %5.1f%%  %6d" % (k[0], k[1], 100.0 * v[0] / tot[0], v[1] / max(v[0], 1), 100.0 * v[2] / max(tot[2], 1), v[3]))


if __name__ == "__main__":
    main()
