#!/usr/bin/env python
"""SASS evidence for profiles/: per hot kernel the opcode histogram of libdibr_b200.so (cuobjdump -sass), the
resource usage (cuobjdump -res-usage) and the full listing gzipped next to it.

  python tools/sass_summary.py profiles/r01_sass
"""
import collections
import gzip
import re
import subprocess
import sys
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "kaolin_b200", "libdibr_b200.so")
HOT = ["setup_count_kernelIf", "scan_kernel", "fill_kernel", "raster_fwd_scatter_kernel", "raster_bwd_kernelIf",
       "soft_mask_bwd_log_kernelIf", "prepare_vertices_bwd_kernelIf",
       "soft_mask_kernelIfLi0", "soft_mask_kernelIfLi1"]


def main():
    prefix = sys.argv[1]
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    usage = {}
    fn = None
    for line in res.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            fn = m.group(1)
        elif fn and "REG:" in line:
            usage[fn] = line.strip()
            fn = None
    blocks, cur = {}, None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            blocks[cur] = []
        elif cur:
            blocks[cur].append(line)
    out = ["# cuobjdump -sass / -res-usage of kaolin_b200/libdibr_b200.so (sm_100a), hot kernels only",
           "# opcode histogram = static instruction counts; full listings: " + os.path.basename(prefix) + "_full.txt.gz", ""]
    full = []
    for key in HOT:
        for name, lines in blocks.items():
            if key not in name:
                continue
            ops = collections.Counter()
            for l in lines:
                m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_]*)(\.[A-Z0-9_.]+)?", l)
                if m:
                    ops[m.group(1) + (m.group(2) or "") if m.group(1) in ("ATOMS", "RED", "ATOMG", "REDUX", "MUFU", "SHFL", "LDG", "STG")
                        else m.group(1)] += 1
            demangled = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
            out.append("== " + demangled[:150])
            out.append("   " + usage.get(name, ""))
            out.append("   instructions: %d" % sum(ops.values()))
            out.append("   " + "  ".join(f"{k}:{v}" for k, v in ops.most_common(40)))
            out.append("")
            full.append("Function : " + name)
            full.extend(lines)
    open(prefix + "_summary.txt", "w").write("\n".join(out) + "\n")
    with gzip.open(prefix + "_full.txt.gz", "wt") as fh:
        fh.write("\n".join(full))
    print("\n".join(out))


if __name__ == "__main__":
    main()
