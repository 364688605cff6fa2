"""Join an ncu SASS source page (csv) with nvdisasm line info: executed warp
instructions and stall samples per CUDA source line of one kernel.

usage: ncu_lines.py <ncu source csv> <nvdisasm --print-line-info output> <kernel substring> [top]
"""
import csv
import re
import sys
from collections import defaultdict


def main():
    src_csv, dis, kern = sys.argv[1], sys.argv[2], sys.argv[3]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    rows = list(csv.reader(open(src_csv)))
    h = next(i for i, r in enumerate(rows) if "Source" in r and "Address" in r)
    ix = {k: j for j, k in enumerate(rows[h])}
    insts = []
    for r in rows[h + 1:]:
        if len(r) < len(rows[h]) or r[ix["Address"]] == "Address":
            continue
        insts.append((r[ix["Source"]].strip(), int(r[ix["Instructions Executed"]]), int(r[ix["# Samples"]]),
                      int(r[ix["Thread Instructions Executed"]])))
    # line info from nvdisasm, in instruction order
    lines = []
    cur = None
    inside = False
    for ln in open(dis):
        if ln.startswith("//---") and ".text." in ln:
            inside = kern in ln
            continue
        if not inside:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', ln)
        if m:
            inl = re.findall(r'inlined at "([^"]+)", line (\d+)', m.group(3))
            cur = (m.group(1).split("/")[-1], int(m.group(2)), tuple((a.split("/")[-1], int(b)) for a, b in inl))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4}\*/", ln):
            lines.append(cur)
    n = min(len(lines), len(insts))
    agg = defaultdict(lambda: [0, 0, 0])
    for k in range(n):
        key = lines[k][:2] if lines[k] else ("?", 0)
        a = agg[key]
        a[0] += insts[k][1]
        a[1] += insts[k][2]
        a[2] += insts[k][3]
    tot = sum(a[0] for a in agg.values()) or 1
    tots = sum(a[1] for a in agg.values()) or 1
    print("instructions (SASS) %d, line-info entries %d, warp inst %d, samples %d" % (len(insts), len(lines), tot, tots))
    for key, a in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
        print("%5.1f%% inst %5.1f%% stall  lanes %4.1f  %s:%d" % (100.0 * a[0] / tot, 100.0 * a[1] / tots,
                                                             a[2] / max(a[0], 1), key[0], key[1]))


main()
