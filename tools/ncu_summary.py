"""Summarise one kernel of an .ncu-rep (run here, no GPU): python tools/ncu_summary.py rep [idx]"""
import csv, subprocess, sys
rep = sys.argv[1]; idx = int(sys.argv[2]) if len(sys.argv) > 2 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units, vals = rows[0], rows[1], rows[2 + idx]
keys = ['Kernel Name', 'gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'sm__warps_active.avg.pct', 'sm__throughput.avg.pct', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'sm__pipe_fp64_cycles_active.avg.pct', 'smsp__issue_active.avg.pct', 'smsp__inst_executed.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared', 'sass__inst_executed_local', '_per_issue_active.ratio',
        'l1tex__data_pipe_lsu_wavefronts', 'lts__t_bytes.sum', 'l1tex__throughput.avg.pct', 'launch__shared_mem_per_block_dynamic',
        'sm__inst_executed_pipe_', 'l1tex__lsu_writeback', 'lts__throughput.avg.pct', 'smsp__inst_executed_op_shared',
        'sm__cycles_elapsed.avg ', 'l1tex__t_bytes_pipe_lsu_mem_global_op_ld.sum ', 'lts__t_sectors_srcunit_tex_op_read.sum ', 'sm__inst_executed_pipe_xu', 'l1tex__data_pipe_lsu_wavefronts_mem_shared']
for h, u, v in zip(hdr, units, vals):
    if any(k in h for k in keys) and 'pct_of_peak_sustained_elapsed' not in h.replace('sm__throughput.avg.pct_of_peak_sustained_elapsed','') :
        print(h, '|', u, '|', v)
