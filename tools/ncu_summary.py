#!/usr/bin/env python
"""Text summary of an `ncu --set full --import-source on` capture of the frame kernel (run here, no GPU needed):
    python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/r02_ncu_full_k_link_frame_t64.txt
Key counters (duration, pipe utilisation, issue slots, DRAM / shared-memory traffic, registers, local memory), the stall
breakdown per issued instruction, the executed-instruction mix by opcode class and the stall samples per code phase."""
import collections
import csv
import io
import re
import subprocess
import sys


def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main(rep):
    rows = page(rep, "raw")
    hdr, units, vals = rows[0], rows[1], rows[2]
    m = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
    print("capture:", rep)
    print("kernel :", m.get("Kernel Name", ("?",))[0], " grid", m.get("Grid Size", ("?",))[0], " block", m.get("Block Size", ("?",))[0])
    keys = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
            "sm__warps_active.avg.per_cycle_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
            "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
            "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
            "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "dram__bytes_read.sum.pct_of_peak_sustained_elapsed", "sass__inst_executed_local_loads", "sass__inst_executed_local_stores",
            "smsp__average_warp_latency_per_inst_issued.ratio"]
    for k in keys:
        if k in m:
            print("  %-78s %s %s" % (k, m[k][0], m[k][1]))
    print("stall cycles per issued instruction:")
    st = [(float(v[0]), h.split("issue_stalled_")[1].split("_per_issue")[0]) for h, v in m.items()
          if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
    for v, name in sorted(st, reverse=True):
        if v > 0.005:
            print("  %-22s %.3f" % (name, v))
    src = page(rep, "source")
    h2 = src[1]
    ix = {h: i for i, h in enumerate(h2)}
    data = src[2:]
    ops = collections.Counter()
    tot_samples = 0
    for r in data:
        mm = re.match(r"\s*(?:@!?U?P[0-9T]+\s+)?([A-Z0-9_]+)", r[1])
        if not mm:
            continue
        ops[mm.group(1)] += int(r[ix["Instructions Executed"]])
        tot_samples += int(r[ix["# Samples"]])
    tot = sum(ops.values())
    print("executed warp instructions by opcode (%d total):" % tot)
    for k, v in ops.most_common(16):
        print("  %-10s %12d  %5.1f %%" % (k, v, 100.0 * v / tot))
    # phases: split the hot loop body at its barriers (BAR / SYNCS) and report samples per segment kind
    seg = collections.Counter()
    for r in data:
        mm = re.match(r"\s*(?:@!?U?P[0-9T]+\s+)?([A-Z0-9_]+)", r[1])
        if not mm:
            continue
        op = mm.group(1)
        kind = ("fp32 (FFMA2/FADD2/FMUL2/FFMA/FMUL)" if op in ("FFMA2", "FADD2", "FMUL2", "FFMA", "FMUL", "FADD") else
                "shared memory (LDS/STS)" if op in ("LDS", "STS") else "barrier / mbarrier" if op in ("BAR", "SYNCS", "WARPSYNC") else
                "MUFU / min-max" if op in ("MUFU", "FMNMX") else "move / integer" if op in ("MOV", "IADD3", "LOP3", "IMAD", "UMOV", "SHF", "LEA", "PLOP3", "ISETP", "SEL", "R2UR", "CS2R", "PRMT", "I2FP") else
                "global memory" if op in ("LDG", "STG", "ATOMG", "RED", "LDC", "LDCU", "UBLKCP", "UBLKPF") else "other")
        seg[kind] += int(r[ix["# Samples"]])
    print("stall samples by the instruction class they were taken at (%d samples):" % tot_samples)
    for k, v in seg.most_common():
        print("  %-40s %6.1f %%" % (k, 100.0 * v / max(tot_samples, 1)))


if __name__ == "__main__":
    main(sys.argv[1])
