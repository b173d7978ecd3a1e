"""Key counters of every kernel in an .ncu-rep: time, instructions, issue rate, pipes, stall reasons, DRAM bytes.
Usage: python tools/ncu_brief.py gpurun_out/x.ncu-rep"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[0]
want = ['gpu__time_duration.sum', 'sm__cycles_elapsed.max', 'smsp__inst_executed.sum', 'sm__inst_executed.avg.per_cycle_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'launch__grid_size',
        'launch__registers_per_thread', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed']
def f(v):
    try: return float(v.replace(',', ''))
    except Exception: return None
for r in rows[2:]:
    print("##", r[hdr.index('Kernel Name')][:90])
    for w in want:
        if w in hdr: print(f"  {w} = {r[hdr.index(w)]} {rows[1][hdr.index(w)]}")
    pipes = [(h, f(r[i])) for i, h in enumerate(hdr) if h.startswith('sm__inst_executed_pipe_') and h.endswith('.avg.pct_of_peak_sustained_active')]
    print("  pipes %:", ", ".join(f"{h.split('pipe_')[1].split('.')[0]} {v:.0f}" for h, v in pipes if v and v > 2))
    st = [(h, f(r[i])) for i, h in enumerate(hdr) if h.startswith('smsp__average_warps_issue_stalled_') and h.endswith('_per_issue_active.ratio')]
    print("  stalls/issue:", ", ".join(f"{h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]} {v:.2f}" for h, v in st if v and v > 0.1))
