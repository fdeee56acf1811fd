"""Summarise an .ncu-rep: key raw metrics + hottest SASS lines. usage: python scripts/ncu_summary.py rep [top]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 20
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
keys = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_bytes.sum', 'lts__t_sectors_srcunit_tex_op_read.sum',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__inst_executed.sum', 'sm__cycles_active.avg', 'launch__registers_per_thread', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct', 'sm__inst_executed_pipe_uniform',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio']
for r in rows[2:]:
    for i, h in enumerate(hdr):
        if h in keys:
            print(f"{h:90s} {r[i]} {rows[1][i]}")
    print('-' * 20)
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h = rows[1]
ismp, iex, isrc = h.index("# Samples"), h.index("Instructions Executed"), h.index("Source")
stall = [i for i, x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
data = [(int(r[ismp] or 0), int(r[iex] or 0), r[isrc], r) for r in rows[2:] if len(r) > iex and (r[ismp] or "0").isdigit()]
print("total samples", sum(d[0] for d in data), "total inst", sum(d[1] for d in data))
for d in sorted(data, key=lambda x: -x[0])[:top]:
    st = sorted(((int(d[3][i] or 0), h[i]) for i in stall), reverse=True)[:2]
    print(f"{d[0]:8d} {d[1]:12d}  {d[2][:90]:90s} {st}")
