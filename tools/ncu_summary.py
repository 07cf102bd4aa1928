#!/usr/bin/env python3
"""Summarise an .ncu-rep (raw + source pages) into a small text file for profiles/.  Runs without a GPU."""
import csv, collections, io, re, subprocess, sys

def page(rep, which, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", which, "--csv", *extra], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))

def main(rep, out_path=None):
    lines = []
    rows = page(rep, "raw")
    hdr, units, data = rows[0], rows[1], rows[2:]
    want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "sm__inst_executed.avg.per_cycle_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "launch__registers_per_thread", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
            "launch__block_size", "launch__grid_size", "launch__shared_mem_per_block_dynamic",
            "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
            "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
            "lts__t_bytes.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__cycles_active.avg", "smsp__inst_executed.sum"]
    want += [h for h in hdr if h.startswith("smsp__average_warp_latency_issue_stalled") or h.startswith("smsp__average_warps_issue_stalled")]
    for w in want:
        if w in hdr:
            i = hdr.index(w)
            lines.append(f"{w} [{units[i]}] = {', '.join(r[i] for r in data)}")
    src = page(rep, "source", ["--print-source", "sass"])
    if len(src) > 2:
        h = src[1]
        iS, iE = h.index("Source"), h.index("Instructions Executed")
        ops = collections.Counter()
        for r in src[2:]:
            if r and r[0] == "Kernel Name":
                break
            if len(r) <= iE:
                continue
            m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_]+)", r[iS].strip())
            ops[m.group(2) if m else r[iS]] += int(r[iE] or 0)
        tot, sad = sum(ops.values()), ops.get("VABSDIFF4", 1)
        lines.append(f"-- executed warp instructions by opcode (first captured launch): total {tot}, VABSDIFF4 share {sad / tot:.4f}")
        alu = ("VABSDIFF4 IADD3 LOP3 SHF PRMT ISETP SEL VIMNMX VIMNMX3 VIADD VIADDMNMX LEA IABS PLOP3 P2R R2P IMNMX POPC FLO BMSK SGXT MOV").split()
        alu_tot = sum(ops[o] for o in alu)
        lines.append(f"-- ALU-pipe class ops {alu_tot}: VABSDIFF4 share of them {sad / alu_tot:.4f}")
        for op, n in ops.most_common(24):
            lines.append(f"   {op:12s} {n:14d} {n / tot * 100:6.2f}%  per 1280 SAD {n / sad * 1280:7.1f}")
    text = "\n".join(lines) + "\n"
    if out_path:
        open(out_path, "w").write(text)
    print(text)

if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else None)
