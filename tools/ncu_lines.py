#!/usr/bin/env python3
"""tools/ncu_lines.py REPORT.ncu-rep [N] -- the N source lines with the most warp-stall samples of an
`ncu --set full --import-source on` report (kernels built with -lineinfo), with the dominant stall reasons.
Reads `ncu -i ... --page source --print-source cuda,sass --csv`."""
import csv
import io
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                         capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    cur, hdr, data, tot = None, None, [], 0
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1].split("/")[-1]
            continue
        if r[0] == "Line No":
            hdr = r
            continue
        if hdr is None or not r[0].isdigit():
            continue
        iS, iI = hdr.index("# Samples"), hdr.index("Instructions Executed")
        try:
            n = int(r[iS])
        except ValueError:
            continue
        stalls = {}
        for k, v in zip(hdr, r):
            if k.startswith("stall_") and "Not Issued" not in k:
                try:
                    stalls[k[6:]] = int(v)
                except ValueError:
                    pass
        tot += n
        data.append((n, cur, int(r[0]), r[1].strip()[:100], int(r[iI] or 0), stalls))
    data.sort(key=lambda x: -x[0])
    print("total samples", tot)
    for n, f, l, s, i, st in data[:top]:
        best = sorted(st.items(), key=lambda kv: -kv[1])[:3]
        print("%6d %5.1f%% %s:%d [%d instr] %s | %s" % (n, 100.0 * n / max(1, tot), f, l, i, s,
                                                        " ".join("%s=%d" % kv for kv in best)))


if __name__ == "__main__":
    main()
