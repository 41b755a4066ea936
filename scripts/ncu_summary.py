#!/usr/bin/env python
"""Condense an .ncu-rep (ncu --set full) into the short text summary committed under profiles/.

    python scripts/ncu_summary.py gpurun_out/prof.ncu-rep [frames_per_launch] > profiles/<name>.txt
"""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.avg", "sm__cycles_active.avg",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed.sum", "smsp__inst_executed.sum", "sm__inst_executed.avg.per_cycle_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_fmaheavy.sum", "sm__inst_executed_pipe_fmalite.sum",
    "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_lsu.sum", "sm__inst_executed_pipe_xu.sum",
    "sm__inst_executed_pipe_uniform.sum", "sm__inst_executed_pipe_cbu.sum",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum",
    "smsp__inst_executed_op_shared_ld.sum", "smsp__inst_executed_op_shared_st.sum",
    "smsp__inst_executed_op_global_ld.sum", "smsp__inst_executed_op_global_st.sum",
    "smsp__inst_executed_op_local_ld.sum", "smsp__inst_executed_op_local_st.sum",
]
STALL_PREFIX = "smsp__average_warps_issue_stalled_"
STALL_SUFFIX = "_per_issue_active.ratio"


def main():
    rep = sys.argv[1]
    frames = float(sys.argv[2]) if len(sys.argv) > 2 else None
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        u = dict(zip(hdr, units))
        print("kernel: %s  grid %s block %s" % (d.get("Kernel Name"), d.get("Grid Size"), d.get("Block Size")))
        for k in KEYS:
            if k in d and d[k] != "":
                print("  %-82s %s %s" % (k, d[k], u.get(k, "")))
        stalls = []
        for k in hdr:
            if k.startswith(STALL_PREFIX) and k.endswith(STALL_SUFFIX) and d[k] not in ("", "0"):
                try:
                    stalls.append((float(d[k].replace(",", "")), k[len(STALL_PREFIX):-len(STALL_SUFFIX)]))
                except ValueError:
                    pass
        stalls.sort(reverse=True)
        print("  warp stall reasons (warps stalled per issue-active cycle):")
        for v, k in stalls[:10]:
            print("    %-40s %.3f" % (k, v))
        if frames:
            def num(k):
                try:
                    return float(d[k].replace(",", ""))
                except Exception:
                    return float("nan")
            def nbytes(k):
                return num(k) * {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1, "Tbyte": 1e12}.get(u.get(k), 1)
            un = u.get("gpu__time_duration.sum", "ns")
            t = num("gpu__time_duration.sum") * {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}.get(un, 1e-9)
            print("  per frame (%.0f frames/launch): %.1f warp-instructions, %.1f shared wavefronts, %.1f DRAM bytes, "
                  "%.1f SM-cycles/frame/SM" % (frames, num("smsp__inst_executed.sum") / frames,
                                                num("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum") / frames,
                                                (nbytes("dram__bytes_read.sum") + nbytes("dram__bytes_write.sum")) / frames,
                                                num("sm__cycles_elapsed.avg") * 148 / frames))
            print("  duration %.3f ms" % (t * 1e3))
            print("  dram traffic per launch: %.0f bytes" % (nbytes("dram__bytes_read.sum") + nbytes("dram__bytes_write.sum")))


if __name__ == "__main__":
    main()
