"""Join an ncu SASS source page (csv) with nvdisasm line info: executed warp-instructions and stall samples per source line.

  ncu -i rep.ncu-rep --page source --csv --kernel-name K > sass.csv
  cuobjdump -xelf all lib.so ; nvdisasm -g file.cubin > dis.txt
  python scripts/ncu_lines.py sass.csv dis.txt <mangled-substring> [top]
"""
import collections
import csv
import re
import sys


def disasm_lines(path, func_sub):
    """offset -> (file, line) for the function whose .text section name contains func_sub"""
    out, cur, active = {}, None, False
    for ln in open(path, errors="replace"):
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
        if m:
            active = func_sub in m.group(1)
            continue
        if re.match(r"\s*\.section\s", ln):
            active = False
        if not active:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);", ln)
        if m:
            out[int(m.group(1), 16)] = (cur, m.group(2))
    return out


def main():
    sass_csv, dis, func = sys.argv[1], sys.argv[2], sys.argv[3]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    lines = disasm_lines(dis, func)
    rows = list(csv.reader(open(sass_csv)))
    h = next(i for i, r in enumerate(rows) if "Instructions Executed" in r)
    hdr = rows[h]
    ia, ii, isamp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples")
    base = None
    agg = collections.defaultdict(lambda: [0, 0])
    tot_i = tot_s = 0
    for r in rows[h + 1:]:
        if len(r) <= ii or not r[ia].startswith("0x"):
            continue
        a = int(r[ia], 16)
        base = a if base is None else base
        n, s = int(r[ii] or 0), int(r[isamp] or 0)
        loc = lines.get(a - base, (None, ""))[0]
        agg[loc][0] += n
        agg[loc][1] += s
        tot_i += n
        tot_s += s
    print("total warp-instructions %d, samples %d, mapped offsets %d" % (tot_i, tot_s, len(lines)))
    for loc, (n, s) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print("%6.2f%% inst  %6.2f%% samples  %s" % (100.0 * n / max(tot_i, 1), 100.0 * s / max(tot_s, 1), loc))


if __name__ == "__main__":
    main()
