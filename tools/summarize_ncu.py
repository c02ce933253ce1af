#!/usr/bin/env python
"""Summarises an ncu launch list (--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
--csv) per kernel, and optionally a `--page raw --csv` export of a full capture of hist_build_k.
  python tools/summarize_ncu.py launches.csv [raw.csv] > profiles/rNN_summary.txt
Writes profiles/hist_build_ncu.json (per-launch DRAM traffic used by bench.py's roofline.traffic)."""
import collections
import csv
import hashlib
import json
import os
import subprocess
import sys

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
HIST_KERNELS = ("hist_span_k", "hist_build_k")          # device-planned level loop / host-planned path


def kernel_source_hash():
    """what bench.py compares before it reports the capture's DRAM traffic: the histogram kernels' source"""
    h = hashlib.sha256()
    for f in ("hist.cu", "common.cuh", "device_util.cuh"):
        h.update(open(os.path.join(ROOT, "ytk-learn_b200", "csrc", f), "rb").read())
    return h.hexdigest()[:16]


def to_us(v, u):
    return {"ns": v / 1e3, "us": v, "ms": v * 1e3, "s": v * 1e6}.get(u, v)


def to_mb(v, u):
    return {"byte": v / 1e6, "Kbyte": v / 1e3, "Mbyte": v, "Gbyte": v * 1e3}.get(u, v)


def main():
    lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
    per = collections.OrderedDict()
    for row in csv.DictReader(lines):
        key = (int(row["ID"]), row["Kernel Name"].split("(")[0].replace("ytk::", "").replace("void ", "").split("<")[0], row["Grid Size"])
        per.setdefault(key, {})[row["Metric Name"]] = (float(row["Metric Value"].replace(",", "")), row["Metric Unit"])
    agg = collections.OrderedDict()
    hist = []
    for (i, name, grid), d in per.items():
        us = to_us(*d["gpu__time_duration.sum"])
        rd = to_mb(*d.get("dram__bytes_read.sum", (0, "byte")))
        wr = to_mb(*d.get("dram__bytes_write.sum", (0, "byte")))
        a = agg.setdefault(name, [0, 0.0, 0.0, 0.0])
        a[0] += 1; a[1] += us; a[2] += rd; a[3] += wr
        if name in HIST_KERNELS:
            hist.append({"id": i, "grid": grid, "us": us, "dram_read_mb": rd, "dram_write_mb": wr})
    tot = sum(a[1] for a in agg.values())
    print("kernel                      launches   total us   share   DRAM read MB  write MB   GB/s")
    for name, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-26s %8d %10.1f  %5.1f%%   %11.1f %9.1f %7.0f" % (name, a[0], a[1], 100 * a[1] / tot, a[2], a[3],
                                                                 (a[2] + a[3]) / a[1] * 1e3 if a[1] else 0))
    print("all kernels: %.1f us" % tot)
    print("\nhist_build_k launches (cold-cache, serialised by ncu):")
    for h in hist[:16]:
        print("  id %4d grid %-14s %8.1f us  read %7.1f MB  write %5.1f MB" % (h["id"], h["grid"], h["us"], h["dram_read_mb"],
                                                                             h["dram_write_mb"]))
    # kernels launched inside ytk_gbdt_make_tree (the timed step); the rest of a bench.py run is the one-off
    # row upload, get_positions and the device-gradient rounds
    in_make = ("hist_", "csr_default_fix_k", "partition_", "plan_", "split_", "leaf_sums_k", "zero_segments_k",
               "grad_absmax_k", "grad_quantize_k", "copy_segments_k", "peer_", "lv_")
    make_tot = sum(a[1] for n, a in agg.items() if n.startswith(in_make))
    hist_us = sum(agg.get(k, [0, 0])[1] for k in HIST_KERNELS)
    print("kernels of make(): %.1f us; histogram-build share of them: %.3f" % (make_tot, hist_us / make_tot if make_tot else 0))
    if hist:
        try:
            commit = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
        except Exception:
            commit = ""
        out = {"kernel": "+".join(k for k in HIST_KERNELS if k in agg), "launches": len(hist),
               "share_of_make_kernel_time": hist_us / make_tot,
               "dram_bytes_per_launch": sum(h["dram_read_mb"] + h["dram_write_mb"] for h in hist) / len(hist) * 1e6,
               "us_per_launch_under_ncu": sum(h["us"] for h in hist) / len(hist),
               "share_of_all_kernel_time_in_the_run": hist_us / tot, "source": os.path.basename(sys.argv[1]),
               "kernel_source_sha256_16": kernel_source_hash(), "summarised_at_commit": commit}
        json.dump(out, open(os.path.join(ROOT, "profiles", "hist_build_ncu.json"), "w"), indent=1)
    if len(sys.argv) > 2:
        rows = list(csv.reader(open(sys.argv[2])))
        hdr = rows[0]
        want = ["Grid Size", "gpu__time_duration.sum", "sm__cycles_elapsed.max", "smsp__inst_executed.sum",
                "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
                "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
                "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
                "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
                "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_atom.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_atom.sum",
                "smsp__inst_executed_op_shared_atom.sum", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
                "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]
        for r in rows[2:]:
            d = dict(zip(hdr, r))
            print("\n--set full capture: %s" % d.get("Kernel Name", "")[:40])
            for k in want:
                if k in d:
                    print("  %-72s %s %s" % (k, d[k], rows[1][hdr.index(k)]))
            st = [(k, float(v.replace(",", ""))) for k, v in d.items() if "issue_stalled" in k and k.endswith("_per_issue_active.ratio")]
            print("  top stall reasons (warps per issue-active cycle): " +
                  ", ".join("%s %.2f" % (k.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), v)
                            for k, v in sorted(st, key=lambda kv: -kv[1])[:5]))


if __name__ == "__main__":
    main()
