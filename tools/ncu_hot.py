"""Hottest SASS lines per kernel from `ncu -i rep --page source --csv` (usage: ncu_hot.py file.ncu-rep [N] [kernel-substring])."""
import csv
import subprocess
import sys


def main(path, top=40, which=""):
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    blocks, cur = [], None
    for r in csv.reader(out.splitlines()):
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "rows": []}
            blocks.append(cur)
        elif cur is not None:
            cur["rows"].append(r)
    for b in blocks:
        if which not in b["name"]:
            continue
        head = b["rows"][0]
        ci = {n: i for i, n in enumerate(head)}
        body = [r for r in b["rows"][1:] if len(r) == len(head)]
        total = sum(int(r[ci["# Samples"]]) for r in body)
        print("====", b["name"][:70], "samples", total, "instructions", len(body))
        stall_cols = [n for n in head if n.startswith("stall_") and "Not Issued" not in n]
        agg = {n: sum(int(r[ci[n]]) for r in body) for n in stall_cols}
        print("  ", ", ".join("%s %.1f%%" % (n[6:], 100.0 * v / total) for n, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
        order = sorted(range(len(body)), key=lambda k: -int(body[k][ci["# Samples"]]))[:top]
        for k in sorted(order):
            r = body[k]
            st = sorted(((int(r[ci[n]]), n[6:]) for n in stall_cols), reverse=True)[:2]
            print("%5d %5.2f%% exec=%-8s %-64s %s" % (k, 100.0 * int(r[ci["# Samples"]]) / total, r[ci["Instructions Executed"]],
                                                     r[ci["Source"]].strip()[:64], " ".join("%s=%d" % (n, v) for v, n in st)))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40, sys.argv[3] if len(sys.argv) > 3 else "")
