"""Summarise an .ncu-rep of the tile kernel: headline metrics, pipe utilisation, stall reasons, and executed
warp-instructions per code region (regions are delimited by branches).  Usage: ncu_summary.py file.ncu-rep"""
import csv, io, re, subprocess, sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, unit, r = rows[0], rows[1], rows[2]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_issued.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__sass_inst_executed_op_shared_ld.sum",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "lts__t_bytes.sum", "launch__registers_per_thread", "launch__grid_size", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.avg.per_cycle_active"]
for i, h in enumerate(hdr):
    if h in want:
        print("%-70s %-8s %s" % (h, unit[i], r[i]))
print("-- stall reasons (warp-cycles per issued instruction share) --")
st = []
for i, h in enumerate(hdr):
    m = re.match(r"smsp__average_warps_issue_stalled_(.*)_per_issue_active\.ratio", h) or \
        re.match(r"smsp__average_warp_latency_issue_stalled_(.*)\.ratio", h)
    if m:
        try: st.append((float(r[i]), m.group(1)))
        except ValueError: pass
for v, n in sorted(st, reverse=True)[:10]:
    print("  %-28s %.3f" % (n, v))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
start = [i for i, x in enumerate(rows) if x and x[0] == "Address"][0]
h2 = rows[start]
body = []
for x in rows[start + 1:]:
    if x and x[0] == "Address": break
    if len(x) >= 6: body.append(x)
iE, iS = h2.index("Instructions Executed"), h2.index("# Samples")
tot = sum(int(x[iE]) for x in body)
print("-- %d SASS rows, %d warp-instructions; regions above 1%% --" % (len(body), tot))
cur = cs = a = 0
for k, x in enumerate(body):
    cur += int(x[iE]); cs += int(x[iS])
    if re.search(r"\b(BRA|BSYNC|EXIT|SYNCS|BAR|CALL|RET)\b", x[1]) or k == len(body) - 1:
        if cur > tot * 0.01:
            print("  rows %4d-%4d (%3d) %10d %5.1f%% samples %6d  %s | %s" % (a, k, k - a + 1, cur, 100.0 * cur / tot, cs, body[a][1].strip()[:34], x[1].strip()[:44]))
        cur = cs = 0; a = k + 1
