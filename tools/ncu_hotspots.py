"""Summarise `ncu -i X.ncu-rep --page source --csv [--kernel-name regex:...]` (first kernel of the file): opcode mix by executed
instructions and by stall samples, shared-memory wavefronts (excessive = bank conflicts), the most sampled SASS lines.

    ncu -i gpurun_out/r02_k4src.ncu-rep --page source --csv > /tmp/src.csv && python tools/ncu_hotspots.py /tmp/src.csv
"""
import collections
import csv
import re
import sys


def main(path, top=16):
    rows = list(csv.reader(open(path)))
    print("kernel:", rows[0][1][:120])
    hdr = rows[1]
    i_s, i_ex = hdr.index("Source"), hdr.index("Instructions Executed")
    i_w, i_we, i_sm = hdr.index("L1 Wavefronts Shared"), hdr.index("L1 Wavefronts Shared Excessive"), hdr.index("# Samples")
    data = []
    for r in rows[2:]:
        if len(r) <= i_we:
            continue
        if r[i_ex] == "Instructions Executed":      # next launch in the same file
            break
        data.append((r[i_s].strip(), int(r[i_ex] or 0), int(r[i_w] or 0), int(r[i_we] or 0), int(r[i_sm] or 0)))
    tot, ts = sum(d[1] for d in data), sum(d[4] for d in data)
    print(f"SASS lines {len(data)}, warp instructions executed {tot}, stall samples {ts}")
    print(f"shared-memory wavefronts {sum(d[2] for d in data)}, of them excessive (bank conflicts) {sum(d[3] for d in data)}")
    by, bys = collections.Counter(), collections.Counter()
    for s, ex, _, _, sm in data:
        op = re.sub(r"^@!?U?P\d+\s+", "", s).split()[0].split(".")[0] if s else "?"
        by[op] += ex
        bys[op] += sm
    print("opcode      executed  samples")
    for op, c in by.most_common(top):
        print(f"  {op:10s} {c / tot * 100:5.1f} %  {bys[op] / max(ts, 1) * 100:5.1f} %")
    print("most excessive shared-memory wavefronts: (SASS, executed, wavefronts, excessive, samples)")
    for d in sorted(data, key=lambda d: -d[3])[:6]:
        print("  ", d)
    print("most sampled lines:")
    for d in sorted(data, key=lambda d: -d[4])[:top]:
        print("  ", d)


if __name__ == "__main__":
    main(sys.argv[1])
