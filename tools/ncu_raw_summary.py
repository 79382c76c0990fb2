"""Per-kernel metric summary from `ncu -i <rep> --page raw --csv` output (the CSV travels, the .ncu-rep is too big):
   python tools/ncu_raw_summary.py gpurun_out/r1_h_full_raw.csv > profiles/r1_h_ncu_summary.txt"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
want = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram rd"), ("dram__bytes_write.sum", "dram wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram % of peak"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem wavefronts"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smem % of peak"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex % of peak"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2 % of peak"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe % active"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots % active"),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma pipe %"),
        ("smsp__inst_executed.sum", "warp instr"), ("launch__registers_per_thread", "regs"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %")]
print("# ncu --set full --clock-control none of `python bench.py --steps 2 --warmup 3 --no-cpu-baseline` (config 2), captured after the same")
print("# command exited 0 without ncu; 41 consecutive launches of the path's kernels (one step).  Every radix pass is listed")
print("# (they differ), the first two weight-gradient launches, and the first instance of every other kernel.")
print("# B200 peaks for reference: HBM 6527.5 GB/s measured (MEASURED_PEAKS.json); shared memory 128 B/clk/SM.")
seen = {}
for r in rows[2:]:
    name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "").replace("tb::", "").replace("mlp::", "")
    k = seen.get(name, 0)
    seen[name] = k + 1
    if not (name.startswith("radix") or (name.startswith("wgrad") and k < 4) or k == 0):
        continue
    grid = r[idx["Grid Size"]] if "Grid Size" in idx else ""
    block = r[idx["Block Size"]] if "Block Size" in idx else ""
    print(f"\n{name}  [instance {k}]  grid {grid} block {block}")
    for m, label in want:
        if m in idx:
            print(f"    {label:24s} {r[idx[m]]:>18s} {units[idx[m]]}")
