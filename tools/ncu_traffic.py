"""Record the DRAM traffic of one kernel from an `ncu --set full` capture in profiles/traffic.json.

usage: ncu_traffic.py <raw csv of `ncu -i x.ncu-rep --page raw --csv`> <kernel substring> <stage> <model> <corpus> <sentences> <source note>

bench.py quotes `roofline.traffic` from this file -- and only while the kernel sources are the
ones the capture was taken from (bench.source_hash()), so the figure can never go stale silently."""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    raw, kern, stage, model, corpus, n, note = sys.argv[1:8]
    rows = list(csv.reader(open(raw)))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    tot, cnt = 0.0, 0
    for r in rows[2:]:
        if kern not in r[ix["Kernel Name"]]:
            continue
        for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            tot += float(r[ix[m]].replace(",", "")) * UNIT[units[ix[m]]]
        cnt += 1
    if cnt == 0:
        raise SystemExit("kernel %s not in %s" % (kern, raw))
    try:
        tab = json.load(open(bench.TRAFFIC_FILE))
    except (OSError, ValueError):
        tab = {}
    tab["%s|%s|%s|%d" % (stage, model, corpus, int(n))] = {
        "traffic": tot / cnt, "kernel": kern, "launches_captured": cnt, "src_sha": bench.source_hash(), "source": note}
    with open(bench.TRAFFIC_FILE, "w") as f:
        json.dump(tab, f, indent=1, sort_keys=True)
    print(tab)


main()
