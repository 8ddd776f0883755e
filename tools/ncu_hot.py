"""Hottest SASS instructions of an ncu report (read here, no GPU needed), with their stall reasons.

    python tools/ncu_hot.py gpurun_out/prof.ncu-rep [top N] [lo hi]      # lo..hi: print that SASS range instead
"""
import csv
import io
import subprocess
import sys


def num(x):
    try:
        return float(x.replace(",", ""))
    except Exception:
        return 0.0


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[1]
    ci = {h: i for i, h in enumerate(hdr)}
    data = [r for r in rows[2:] if len(r) > ci["# Samples"]]
    stall = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    tot = sum(num(r[ci["# Samples"]]) for r in data)

    def show(i, r):
        st = sorted(((num(r[ci[h]]), h[6:]) for h in stall if num(r[ci[h]]) > 0), reverse=True)[:3]
        print("%5d %5.2f%% %-70s %s" % (i, 100 * num(r[ci["# Samples"]]) / tot, r[ci["Source"]].strip()[:70],
                                        " ".join("%s:%d" % (n, v) for v, n in st)))

    if len(sys.argv) > 4:
        lo, hi = int(sys.argv[3]), int(sys.argv[4])
        for i in range(lo, hi + 1):
            show(i, data[i])
        return
    order = sorted(range(len(data)), key=lambda i: -num(data[i][ci["# Samples"]]))[:top]
    for i in sorted(order):
        show(i, data[i])


if __name__ == "__main__":
    main()
