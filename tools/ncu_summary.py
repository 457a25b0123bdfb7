#!/usr/bin/env python
"""Per-kernel summary of an `ncu --set full` report (read here, no GPU): duration, DRAM bytes, throughputs,
occupancy, top stall reasons.  Writes a CSV for profiles/ and optionally merges DRAM traffic per launch into
profiles/traffic.json (read by bench.py's roofline leg).

  python tools/ncu_summary.py gpurun_out/prof_c2.ncu-rep profiles/r01_c2_ncu_summary.csv [--traffic c2]
"""
import csv
import json
import os
import subprocess
import sys

KEYS = [("gpu__time_duration.sum", "duration"), ("dram__bytes_read.sum", "dram_read"),
        ("dram__bytes_write.sum", "dram_write"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_active_pct"),
        ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "fma_pipe_pct"),
        ("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "alu_pipe_pct"),
        ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "fp64_pipe_pct"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved_occupancy_pct"),
        ("launch__registers_per_thread", "registers"), ("sm__inst_executed.sum", "warp_instructions"),
        ("smsp__thread_inst_executed_per_inst_executed.ratio", "active_threads_per_warp")]
NAMES = {"setup_count_kernel": "setup_count", "fill_kernel": "fill", "scan_kernel": "scan",
         "raster_fwd_scatter_kernel": "raster_fwd", "raster_fwd_kernel": "raster_fwd", "raster_bwd_kernel": "raster_bwd",
         "soft_mask_kernel<float, 0": "soft_fwd", "soft_mask_kernel<float, 1": "soft_bwd_recompute",
         "soft_mask_bwd_log_kernel": "soft_bwd", "soft_mask_overflow_kernel": "soft_bwd_overflow",
         "soft_mask_kernel<double, 0": "soft_fwd_f64", "soft_mask_kernel<double, 1": "soft_bwd_f64"}
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3,
        "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}


def main():
    rep, out = sys.argv[1], sys.argv[2]
    if rep.endswith(".csv") or rep.endswith(".csv.gz"):      # a raw page exported on the GPU box (tools/prof_all.sh)
        import gzip
        raw = (gzip.open(rep, "rt") if rep.endswith(".gz") else open(rep)).read()
    else:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    lines, traffic = [], {}
    for r in rows[2:]:
        name = r[idx["Kernel Name"]]
        short = next((v for k, v in NAMES.items() if k in name), name.split("(")[0][-40:])
        rec = {"kernel": short}
        for key, lab in KEYS:
            if key in idx and r[idx[key]] != "":
                v = float(r[idx[key]].replace(",", "")) * UNIT.get(units[idx[key]], 1.0)
                rec[lab] = v
        stalls = {h.split("issue_stalled_")[1].split("_per_issue")[0]: float(r[i]) for h, i in idx.items()
                  if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("per_issue_active.ratio") and r[i]}
        rec["top_stalls"] = " ".join(f"{k}={v:.2f}" for k, v in sorted(stalls.items(), key=lambda kv: -kv[1])[:4])
        lines.append(rec)
        if "dram_read" in rec:
            traffic[short] = rec["dram_read"] + rec["dram_write"]
    cols = ["kernel"] + [lab for _, lab in KEYS] + ["top_stalls"]
    with open(out, "w") as fh:
        fh.write("# ncu --set full --clock-control none, one launch per kernel; duration in us, dram_* in bytes\n")
        w = csv.writer(fh)
        w.writerow(cols)
        for rec in lines:
            w.writerow([("%.4g" % rec[c]) if isinstance(rec.get(c), float) else rec.get(c, "") for c in cols])
    print(open(out).read())
    if "--traffic" in sys.argv:
        wl = sys.argv[sys.argv.index("--traffic") + 1]
        path = os.path.join(os.path.dirname(out) or ".", "traffic.json")
        allt = json.load(open(path)) if os.path.exists(path) else {}
        allt[wl] = traffic
        json.dump(allt, open(path, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
