#!/usr/bin/env python
"""Aggregate an ncu SASS source page by CUDA source line.

  ncu -i X.ncu-rep --page source --csv --print-source sass > sass.csv
  cuobjdump -xelf all libdibr_b200.so ; nvdisasm -g -c *.cubin > dis.txt
  python tools/ncu_lines.py sass.csv dis.txt [kernel-substring]

Instructions appear in the same order in both listings; inlined code is attributed to the innermost line.
"""
import collections
import csv
import re
import sys


def parse_dis(path):
    funcs, cur, line = {}, None, None
    for l in open(path):
        m = re.match(r"\.text\.(\S+):", l)
        if m:
            cur = m.group(1); funcs[cur] = []; line = None
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            line = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        if cur and re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
            funcs[cur].append(line)
    return funcs


_SRC = {}


def src_text(ln):
    """Source text of (file, line) for files under kaolin_b200/csrc."""
    if not ln:
        return ""
    import os
    f, n = ln
    if f not in _SRC:
        root = os.environ.get("DIBR_SRC_ROOT") or os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        path = os.path.join(root, "kaolin_b200", "csrc", f)      # DIBR_SRC_ROOT: the tree the profiled .so was built from
        _SRC[f] = open(path).read().split("\n") if os.path.exists(path) else []
    return _SRC[f][n - 1].strip()[:90] if 0 < n <= len(_SRC[f]) else ""


def main():
    sass, dis = sys.argv[1], sys.argv[2]
    filt = sys.argv[3] if len(sys.argv) > 3 else ""
    funcs = parse_dis(dis)
    import subprocess
    demangled = {}
    for k in funcs:
        demangled[subprocess.run(["c++filt", k], capture_output=True, text=True).stdout.strip()] = k
    rows = list(csv.reader(open(sass)))
    i = 0
    while i < len(rows):
        if rows[i] and rows[i][0] == "Kernel Name":
            name = rows[i][1]
            hdr = rows[i + 1]
            j = i + 2
            body = []
            while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
                if rows[j]:
                    body.append(rows[j])
                j += 1
            i = j
            if filt.replace(" ", "") not in name.replace("(int)", "").replace(" ", ""):
                continue
            norm = lambda t: t.replace("(int)", "").replace("(bool)1", "true").replace("(bool)0", "false").replace(" ", "").split("(dibr::")[0]
            key = next((v for d, v in demangled.items() if norm(d) == norm(name)), None)
            lines = funcs.get(key, [])
            ci = hdr.index("Instructions Executed"); cs = hdr.index("# Samples")
            agg = collections.defaultdict(lambda: [0, 0])
            tot = 0
            for n, r in enumerate(body):
                ln = lines[n] if n < len(lines) else None
                v = int(r[ci] or 0); agg[ln][0] += v; agg[ln][1] += int(r[cs] or 0); tot += v
            print(f"== {name[:80]}  instr={tot} sass_rows={len(body)} dis_rows={len(lines)}")
            top = int(sys.argv[4]) if len(sys.argv) > 4 else 28
            for ln, (v, s) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
                print(f"   {str(ln):34s} {v:11d} {100.0 * v / max(tot, 1):5.1f}%  smp {s:5d}  {src_text(ln)}")
        else:
            i += 1


if __name__ == "__main__":
    main()
