"""Cut launch number k (0-based) out of a multi-launch ncu export (prefix_details.txt, prefix_raw.csv, prefix_sass.csv.gz)
into outprefix_*, so that tools/ncu_summary.py can summarise it: python tools/ncu_pick.py <prefix> <k> <outprefix>"""
import csv, gzip, re, sys
pre, k, out = sys.argv[1], int(sys.argv[2]), sys.argv[3]
# details: sections start at a non-indented-by-4 kernel line "  <name>(...) (grid)x(block), Context"
lines = open(pre + "_details.txt").read().split("\n")
starts = [i for i, l in enumerate(lines) if re.match(r"^  \S.*Context \d+, Stream", l)]
starts.append(len(lines))
open(out + "_details.txt", "w").write("\n".join(lines[starts[k]:starts[k + 1]]) + "\n")
rows = list(csv.reader(open(pre + "_raw.csv")))
with open(out + "_raw.csv", "w", newline="") as fh:
    w = csv.writer(fh); w.writerow(rows[0]); w.writerow(rows[1]); w.writerow(rows[2 + k])
blocks, cur = [], None
with gzip.open(pre + "_sass.csv.gz", "rt") as fh:
    for r in csv.reader(fh):
        if r and r[0] == "Address":
            cur = [r]; blocks.append(cur)
        elif cur is not None:
            cur.append(r)
per = max(1, len(blocks) // (len(rows) - 2))  # ncu prints the source page of a launch once per view
with gzip.open(out + "_sass.csv.gz", "wt", newline="") as fh:
    csv.writer(fh).writerows(blocks[k * per])
print("launch %d of %d: %d SASS rows" % (k, len(rows) - 2, len(blocks[k * per]) - 1))
