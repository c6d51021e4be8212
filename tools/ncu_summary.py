#!/usr/bin/env python
"""Summarise one `ncu --set full --import-source on` capture of a kernel into the text file kept
under profiles/ (the .ncu-rep itself stays in gpurun_out/, which is scratch).

    python tools/ncu_summary.py gpurun_out/k_edges.ncu-rep [kernel-name-substring] > profiles/r01_k_edges_ncu_summary.txt

With a kernel name the first launch whose name contains it is summarised (a capture of the whole step
holds one launch of every kernel); without, the last launch of the report.

First block: `metric value` lines (bench.py reads dram__bytes_read.sum / dram__bytes_write.sum, in
Mbyte, from it), then the warp-stall reasons per issue-active warp, then the instruction count and
stall-sample share per CUDA source line (needs -lineinfo at compile time)."""
import collections
import csv
import io
import subprocess
import sys

METRICS = ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
           "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
           "smsp__thread_inst_executed_per_inst_executed.ratio", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
           "lts__t_bytes.sum", "lts__t_sectors_op_read.sum", "lts__t_sectors_op_write.sum", "l1tex__t_bytes.sum",
           "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_ltcfabric.sum", "nvlrx__bytes.sum", "nvltx__bytes.sum",
           "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
           "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"]


def ncu(rep, *args):
    return subprocess.run(["ncu", "-i", rep, "--csv", *args], capture_output=True, text=True, check=True).stdout


def main():
    rep = sys.argv[1]
    kern = sys.argv[2] if len(sys.argv) > 2 else None
    rows = list(csv.reader(io.StringIO(ncu(rep, "--page", "raw"))))
    hdr, units, vals = rows[0], rows[1], rows[-1]
    sel = []
    if kern:
        ki = hdr.index("Kernel Name")
        match = [r for r in rows[2:] if len(r) > ki and kern in r[ki]]
        if not match:
            raise SystemExit("no launch of a kernel named *%s* in %s" % (kern, rep))
        vals = match[0]
        sel = ["-k", "regex:" + kern, "-c", "1"]
    print("# %s" % vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "# kernel")
    for m in METRICS:
        if m in hdr:
            print(m, vals[hdr.index(m)], units[hdr.index(m)])
    print("# warp stall reasons, warps per issue-active cycle (> 0.3)")
    stalls = []
    for h, v in zip(hdr, vals):
        if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio"):
            try:
                if float(v) > 0.3:
                    stalls.append((float(v), h.split("issue_stalled_")[1].split("_per_issue")[0]))
            except ValueError:
                pass
    for v, name in sorted(stalls, reverse=True):
        print("%.3f %s" % (v, name))

    rows = list(csv.reader(io.StringIO(ncu(rep, *sel, "--page", "source", "--print-source", "cuda,sass"))))
    agg = collections.defaultdict(lambda: [0, 0, 0])
    fname, have_hdr, tot_i, tot_s = None, False, 0, 0
    for r in rows:
        if len(r) == 2 and r[0] in ("File Path", "File Name"):
            fname = r[1].split("/")[-1]
            continue
        if r and r[0] == "Line No":
            have_hdr = True
            continue
        if not have_hdr or len(r) < 9 or r[2] != "-":
            continue
        try:
            n, s, th = int(r[7]), int(r[4]), int(r[8])
        except ValueError:
            continue
        a = agg[(fname, r[0], r[1].strip())]
        a[0] += n
        a[1] += s
        a[2] += th
        tot_i += n
        tot_s += s
    print("# per CUDA source line: share of executed warp instructions, share of stall samples, threads per instruction")
    print("# totals: %d instructions (lines shared by inlined copies are counted once per copy), %d samples" % (tot_i, tot_s))
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:40]:
        print("%s:%s inst %5.1f%% samp %5.1f%% thr %4.1f  %s" % (k[0], k[1], 100.0 * v[0] / max(tot_i, 1), 100.0 * v[1] / max(tot_s, 1),
                                                             v[2] / max(v[0], 1), k[2][:120]))


if __name__ == "__main__":
    main()
