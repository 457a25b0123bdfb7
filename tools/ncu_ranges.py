#!/usr/bin/env python
"""Aggregate an ncu SASS source page by ranges of source lines (phases of a kernel).

  python tools/ncu_ranges.py sass.csv dis.txt <kernel-substring> file:lo-hi[:label] ...
"""
import collections, csv, subprocess, sys
sys.path.insert(0, __import__("os").path.dirname(__file__))
from ncu_lines import parse_dis

def main():
    sass, dis, filt = sys.argv[1:4]
    ranges = []
    for a in sys.argv[4:]:
        parts = a.split(":")
        lo, hi = (int(x) for x in parts[1].split("-"))
        ranges.append((parts[0], lo, hi, parts[2] if len(parts) > 2 else a))
    funcs = parse_dis(dis)
    dem = {subprocess.run(["c++filt", k], capture_output=True, text=True).stdout.strip(): k for k in funcs}
    rows = list(csv.reader(open(sass)))
    i = 0
    while i < len(rows):
        if rows[i] and rows[i][0] == "Kernel Name":
            name = rows[i][1]; hdr = rows[i + 1]; j = i + 2; body = []
            while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
                if rows[j]: body.append(rows[j])
                j += 1
            i = j
            if filt.replace(" ", "") not in name.replace("(int)", "").replace(" ", ""): continue
            norm = lambda t: t.replace("(int)", "").replace(" ", "").split("(dibr::")[0]
            key = next((v for d, v in dem.items() if norm(d) == norm(name)), None)
            lines = funcs.get(key, [])
            ci = hdr.index("Instructions Executed"); cs = hdr.index("# Samples")
            agg = collections.OrderedDict((r[3], [0, 0]) for r in ranges); agg["other"] = [0, 0]
            oth = collections.defaultdict(int)
            tot = 0; ts = 0
            for n, r in enumerate(body):
                ln = lines[n] if n < len(lines) else None
                v = int(r[ci] or 0); s = int(r[cs] or 0); tot += v; ts += s
                lab = "other"
                if ln:
                    for f, lo, hi, l in ranges:
                        if ln[0] == f and lo <= ln[1] <= hi: lab = l; break
                agg[lab][0] += v; agg[lab][1] += s
                if lab == "other": oth[ln] += v
            print(f"== {name[:70]} instr={tot} samples={ts}")
            for l, (v, s) in agg.items():
                print(f"   {l:40s} {v:11d} {100.0*v/max(tot,1):5.1f}%   samples {100.0*s/max(ts,1):5.1f}%")
            print("   other top:", sorted(oth.items(), key=lambda kv: -kv[1])[:8])
            return
        else:
            i += 1
main()
