"""Per source line of one kernel: share of the executed warp instructions, active lanes per instruction, share of the stall samples.

  ncu -i rep.ncu-rep --page source --csv --kernel-name regex:K > sass.csv
  nvdisasm -g trace.cubin > dis.txt           (the cubin of the SAME build: `nvcc -cubin` of the source file with the build's flags)
  python scripts/ncu_lines_lanes.py sass.csv dis.txt <mangled-substring> [top]
"""
import collections
import csv
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ncu_lines as NL  # noqa: E402


def main():
    sass_csv, dis, func = sys.argv[1], sys.argv[2], sys.argv[3]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 30
    lines = NL.disasm_lines(dis, func)
    rows = list(csv.reader(open(sass_csv)))
    h = next(i for i, r in enumerate(rows) if "Instructions Executed" in r)
    hdr = rows[h]
    ia, ii, it, isamp = (hdr.index(k) for k in ("Address", "Instructions Executed", "Thread Instructions Executed", "# Samples"))
    agg = collections.defaultdict(lambda: [0, 0, 0])
    base, ti, tt, ts = None, 0, 0, 0
    seen = set()
    for r in rows[h + 1:]:
        if len(r) <= it or not r[ia].startswith("0x"):
            continue
        a = int(r[ia], 16)
        if a in seen:      # the page lists the function once per view
            continue
        seen.add(a)
        base = a if base is None else base
        n, t, s = int(r[ii] or 0), int(r[it] or 0), int(r[isamp] or 0)
        loc = lines.get(a - base, (None, ""))[0]
        agg[loc][0] += n
        agg[loc][1] += t
        agg[loc][2] += s
        ti += n
        tt += t
        ts += s
    print("%.1f M warp instructions, %.2f lanes per instruction, %d stall samples, %d mapped offsets" % (ti / 1e6, tt / max(ti, 1), ts, len(lines)))
    print("| instructions | lanes | stall samples | file:line |")
    print("|---|---|---|---|")
    for loc, (n, t, s) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        where = "%s:%d" % loc if loc else "?"
        print("| %.2f %% | %.1f | %.2f %% | %s |" % (100.0 * n / ti, t / max(n, 1), 100.0 * s / max(ts, 1), where))


if __name__ == "__main__":
    main()
