"""Per-source-line instruction / sample shares of one kernel from an ncu report captured with --import-source on.
usage: python tools/ncu_lines.py REPORT.ncu-rep [top]"""
import csv, subprocess, sys

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
cur, data = None, []
for r in csv.reader(out.splitlines()):
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif len(r) > 8 and r[0] not in ("", "Line No"):
        try:
            data.append((int(r[7]), int(r[6]), cur, r[0], r[1][:105]))
        except ValueError:
            pass
tot, ts = sum(d[0] for d in data), sum(d[1] for d in data)
print("total warp instructions", tot, "samples", ts)
for d in sorted(data, reverse=True)[: int(sys.argv[2]) if len(sys.argv) > 2 else 40]:
    print("%5.1f%% inst %5.1f%% smp %s:%s | %s" % (100 * d[0] / tot, 100 * d[1] / ts, d[2], d[3], d[4]))
