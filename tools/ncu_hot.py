"""Hottest SASS lines of a kernel from `ncu -i rep --page source --csv` (usage: ncu_hot.py file.ncu-rep [N])."""
import csv
import subprocess
import sys


def main(path, top=40):
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    head = rows[1]
    ci = {n: i for i, n in enumerate(head)}
    body = [r for r in rows[2:] if len(r) == len(head)]
    total = sum(int(r[ci["# Samples"]]) for r in body)
    print("total samples", total, "instructions", len(body))
    stall_cols = [n for n in head if n.startswith("stall_") and "Not Issued" not in n]
    order = sorted(range(len(body)), key=lambda k: -int(body[k][ci["# Samples"]]))[:top]
    for k in sorted(order):
        r = body[k]
        st = sorted(((int(r[ci[n]]), n[6:]) for n in stall_cols), reverse=True)[:2]
        print("%5d %6.2f%% exec=%-8s %-60s %s" % (k, 100.0 * int(r[ci["# Samples"]]) / total, r[ci["Instructions Executed"]],
                                                 r[ci["Source"]].strip()[:60], " ".join("%s=%d" % (n, v) for v, n in st)))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
