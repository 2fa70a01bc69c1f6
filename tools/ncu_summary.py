"""Summarise an .ncu-rep (one kernel): key raw metrics + executed-instruction blocks from the source page."""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__inst_executed_pipe_alu.avg.pct", "sm__inst_executed_pipe_fp64.avg.pct",
        "sm__inst_executed_pipe_fma.avg.pct", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct", "smsp__inst_executed.sum ", "launch__registers_per_thread ",
        "sm__cycles_elapsed.max ", "smsp__average_warps_issue_stalled", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__warps_active.avg.pct",
        "launch__grid_size", "launch__block_size", "lts__t_bytes.sum ", "sm__throughput.avg.pct", "dram__throughput.avg.pct", "smsp__thread_inst_executed_per_inst_executed.ratio"]
for h, u, v in zip(hdr, units, vals):
    hh = h + " "
    if any(w in hh for w in want) and ".max." not in h and ".min." not in h and ".sum.p" not in h:
        print("%-90s %-12s %s" % (h, u, v))
if len(sys.argv) > 2 and sys.argv[2] == "src":
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    hdr = rows[1]; data = rows[2:]
    ia = hdr.index("Instructions Executed"); isamp = hdr.index("# Samples")
    tot = sum(int(r[ia]) for r in data); ts = sum(int(r[isamp]) for r in data)
    print("source page: total warp instructions", tot, "stall samples", ts)
    blocks = []; cur = None
    for i, r in enumerate(data):
        n = int(r[ia]); sm = int(r[isamp])
        if cur and abs(n - cur['n']) <= 0.02 * max(n, cur['n'], 1):
            cur['cnt'] += 1; cur['tot'] += n; cur['samp'] += sm; cur['end'] = i
        else:
            cur = {'start': i, 'end': i, 'n': n, 'cnt': 1, 'tot': n, 'samp': sm}; blocks.append(cur)
    for b in blocks:
        if b['tot'] > tot * 0.01 or b['samp'] > ts * 0.02:
            print("  sass lines %4d-%4d executed %8d x %3d lines = %5.1f%% of instructions, %5d samples  [%s]" % (b['start'], b['end'], b['n'], b['cnt'], 100 * b['tot'] / tot, b['samp'], data[b['start']][1].strip()[:40]))
    top = sorted(range(len(data)), key=lambda i: -int(data[i][isamp]))[:12]
    print("  top stall lines:")
    for i in sorted(top): print("   ", i, data[i][ia], data[i][isamp], data[i][1].strip()[:90])
