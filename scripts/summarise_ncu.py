"""Condense an `ncu --page raw --csv` export into the one-row-per-kernel summaries kept under profiles/.

usage: python scripts/summarise_ncu.py gpurun_out/raw.csv profiles/out.csv [kernel-substring]
Keeps the last launch of every distinct (kernel, grid, instruction count) / duration triple, i.e. in practice the launches of the final repetition.""" 
import csv
import sys

COLS = [("time", "gpu__time_duration.sum"), ("grid", "launch__grid_size"), ("block", "launch__block_size"),
        ("regs", "launch__registers_per_thread"), ("dram_rd", "dram__bytes_read.sum"),
        ("dram_wr", "dram__bytes_write.sum"), ("dram_pct", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
        ("dram_bytes_per_s", "dram__bytes.sum.per_second"),
        ("sm_pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
        ("occ_pct", "sm__warps_active.avg.pct_of_peak_sustained_active"),
        ("fma_pct", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
        ("alu_pct", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
        ("xu_pct", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
        ("issue_pct", "sm__inst_issued.avg.pct_of_peak_sustained_active"), ("inst", "smsp__inst_executed.sum"),
        ("l2_hit", "lts__t_sector_hit_rate.pct"),
        ("tensor_pct", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
        ("tensor_subpipe_hmma_pct", "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active"),
        ("tensor_mem_pct", "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed"),
        ("smem_bytes", "launch__shared_mem_per_block_dynamic"),
        ("nvlink_rx", "nvlrx__bytes.sum"), ("nvlink_tx", "nvltx__bytes.sum")]


def main():
    src, dst = sys.argv[1], sys.argv[2]
    needle = sys.argv[3] if len(sys.argv) > 3 else ""
    rows = list(csv.reader(open(src, newline="")))
    header, units, body = rows[0], rows[1], rows[2:]
    idx = {name: i for i, name in enumerate(header)}
    kcol = idx["Kernel Name"]
    keep = {}
    for r in body:
        if needle in r[kcol]:
            keep[(r[kcol], r[idx["launch__grid_size"]], r[idx["gpu__time_duration.sum"]])] = r
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel"] + [c for c, m in COLS if m in idx])
        for (k, _, _), r in keep.items():
            w.writerow([k] + [f"{r[idx[m]]} {units[idx[m]]}".strip() for c, m in COLS if m in idx])


if __name__ == "__main__":
    main()
