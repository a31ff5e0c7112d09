"""Per-source-line executed warp instructions from an .ncu-rep (needs -lineinfo + --import-source on):
    python profiles/hot_lines.py gpurun_out/prof.ncu-rep [top_n]
"""
import csv
import subprocess
import sys


def main(rep, top=40):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    fname, ie, line, src = None, None, None, None
    agg = {}
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            fname = r[1].split("/")[-1]
            continue
        if r[0] == "Line No":
            ie = r.index("Instructions Executed") if "Instructions Executed" in r else None
            continue
        if r[0] not in ("", "Function Name", "Kernel Name") and r[0].isdigit():
            line, src = int(r[0]), r[1]
            continue
        if r[0] == "" and ie is not None and len(r) > ie and line is not None:
            try:
                v = int(r[ie])
            except ValueError:
                continue
            k = (fname, line)
            agg.setdefault(k, [0, src])
            agg[k][0] += v
    tot = sum(v[0] for v in agg.values())
    print("total warp instructions executed:", tot)
    for (f, l), (v, s) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print("%10d %5.1f%%  %s:%d  %s" % (v, 100.0 * v / tot, f, l, s[:100]))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
