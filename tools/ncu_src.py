#!/usr/bin/env python
"""Per-instruction stall samples of one kernel from an `ncu --page source --csv --print-source sass` export.
usage: ncu_src.py export.csv kernel_substring [min_samples]"""
import csv
import sys


def main():
    path, key = sys.argv[1], sys.argv[2]
    thr = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    rows = list(csv.reader(open(path)))
    i = 0
    while i < len(rows):
        if rows[i] and rows[i][0] == "Kernel Name" and key in rows[i][1]:
            hdr = rows[i + 1]
            j = i + 2
            body = []
            while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
                body.append(rows[j])
                j += 1
            col = {n: k for k, n in enumerate(hdr)}
            stalls = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
            tot = sum(int(r[col["# Samples"]]) for r in body if len(r) > col["# Samples"])
            print("kernel", rows[i][1], "instructions", len(body), "samples", tot)
            acc = 0
            for r in body:
                if len(r) <= col["# Samples"]:
                    continue
                n = int(r[col["# Samples"]])
                acc += n
                if n >= thr:
                    top = sorted(((int(r[col[s]]), s[6:]) for s in stalls), reverse=True)[:3]
                    print("%6d %5.1f%% cum %5.1f%%  x%-9s %-58s %s" % (
                        n, 100.0 * n / tot, 100.0 * acc / tot, r[col["Instructions Executed"]], r[col["Source"]].strip()[:58],
                        " ".join("%s=%d" % (s, v) for v, s in top if v)))
            return
        i += 1


if __name__ == "__main__":
    main()
