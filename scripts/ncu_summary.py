"""Summarise an .ncu-rep (first kernel): key metrics + per-source-line hot spots. usage: ncu_summary.py REP [top]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; top = sys.argv[2] if len(sys.argv) > 2 else "30"
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, u, v = rows[0], rows[1], rows[2]
keys = ['Kernel Name', 'gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
        'launch__shared_mem_per_block_dynamic', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum', 'l1tex__t_sector_hit_rate.pct',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__inst_executed_op_shared_ld.sum', 'smsp__inst_executed_op_shared_st.sum', 'smsp__inst_executed_op_local_ld.sum']
for i, n in enumerate(h):
    if n in keys or ('issue_stalled' in n and 'per_issue_active' in n and float(v[i] or 0) > 0.05):
        print(f"{n} [{u[i]}] = {v[i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
open("/tmp/_src.csv", "w").write(src)
print(subprocess.run([sys.executable, __file__.replace("ncu_summary.py", "ncu_lines.py"), "/tmp/_src.csv", top], capture_output=True, text=True).stdout)
