"""Key metrics + top stall lines from an .ncu-rep (uses `ncu -i`)."""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]; idx = {h: i for i, h in enumerate(hdr)}
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_red.sum", "lts__t_sectors_op_red.sum", "smsp__inst_executed_op_shfl.sum"]
for r in rows[2:]:
    print("kernel:", r[idx["Kernel Name"]][:80])
    for w in want:
        if w in idx: print("  %-70s %s %s" % (w, r[idx[w]], rows[1][idx[w]]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]; si = hdr.index("Source"); ws = hdr.index("Warp Stall Sampling (All Samples)"); ie = hdr.index("Instructions Executed")
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
data = []; agg = {}
for k, r in enumerate(rows[2:]):
    if len(r) < len(hdr): break
    try: v = float(r[ws])
    except ValueError: continue
    for i, h in stall_cols:
        agg[h] = agg.get(h, 0) + float(r[i])
    st = sorted(((float(r[i]), h) for i, h in stall_cols), reverse=True)[:2]
    data.append((k, v, r[si].strip()[:60], st, r[ie]))
tot = sum(d[1] for d in data)
print("stall totals:", sorted(((int(v), k[6:]) for k, v in agg.items() if v > 0), reverse=True)[:8])
op = {}
for d in data:
    o = d[2].split()[0] if not d[2].startswith("@") else d[2].split()[1]
    op[o.split(".")[0]] = op.get(o.split(".")[0], 0) + float(d[4] or 0)
print("instr mix:", sorted(((int(v), k) for k, v in op.items()), reverse=True)[:14])
for d in sorted(data, key=lambda d: -d[1])[: int(sys.argv[2]) if len(sys.argv) > 2 else 20]:
    print("%5.1f%% #%4d %-60s %s" % (100 * d[1] / tot, d[0], d[2], [(h[6:], int(x)) for x, h in d[3] if x > 0]))
