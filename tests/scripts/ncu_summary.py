"""Summarise an ncu report (.ncu-rep) into the text kept under profiles/: per captured launch the duration, DRAM bytes,
issue utilisation, registers/smem and the warp-stall breakdown (smsp__average_warps_issue_stalled_*_per_issue_active).
usage: python tests/scripts/ncu_summary.py report.ncu-rep [every_nth] > profiles/xxx.txt"""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "smsp__cycles_active.avg", "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_fma.sum",
        "sm__inst_executed_pipe_fmaheavy.sum", "sm__inst_executed_pipe_tensor.sum", "sm__inst_executed_pipe_lsu.sum",
        "sm__inst_executed_pipe_uniform.sum", "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_op_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor_op_hmma.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]


def main():
    rep = sys.argv[1]
    nth = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    stall = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio")]
    print(f"# {rep}: {len(data)} captured launches, every {nth}. one shown (ncu --set full --clock-control none; cold-cache, serialised)")
    for k, r in enumerate(data):
        if k % nth != nth - 1:
            continue
        print(f"\n== launch {k}: {r[idx['Kernel Name']]}  grid={r[idx['Grid Size']]} block={r[idx['Block Size']]}")
        for w in WANT:
            if w in idx and r[idx[w]] not in ("", "n/a"):
                print(f"  {w:75s} {r[idx[w]]} {units[idx[w]]}")
        st = sorted(((float(r[idx[h]].replace(",", "")), h) for h in stall if r[idx[h]] not in ("", "n/a")), reverse=True)
        print("  warp stalls (warps per issue-active cycle):")
        for v, h in st:
            if v >= 0.01:
                print(f"    {v:6.2f}  {h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')}")


if __name__ == "__main__":
    main()
