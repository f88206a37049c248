"""Summarise an .ncu-rep (raw page) into the handful of counters the roofline discussion needs.
usage: python profiles/ncu_summary.py gpurun_out/prof.ncu-rep"""
import csv, subprocess, sys, io

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'sm__inst_executed_pipe_fma.sum', 'sm__inst_executed_pipe_alu.sum', 'sm__inst_executed_pipe_lsu.sum', 'sm__inst_executed_pipe_fp64.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'smsp__sass_inst_executed_op_local_ld.sum',
        'smsp__sass_inst_executed_op_local_st.sum', 'smsp__sass_inst_executed_op_shared_ld.sum', 'smsp__sass_inst_executed_op_shared_st.sum',
        'smsp__sass_inst_executed_op_global_ld.sum', 'smsp__cycles_active.avg']

def main():
    rep = sys.argv[1]
    out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h, units = rows[0], rows[1]
    for r in rows[2:]:
        print('== kernel:', r[h.index('Kernel Name')] if 'Kernel Name' in h else '?')
        for i, n in enumerate(h):
            pipe = n.startswith('sm__inst_executed_pipe_') and (n.endswith('.sum') or n.endswith('.avg.pct_of_peak_sustained_active'))
            if n in KEYS or pipe or (n.startswith('smsp__average_warps_issue_stalled') and n.endswith('per_issue_active.ratio')):
                print(f'{n} = {r[i]} {units[i]}')

if __name__ == '__main__':
    main()
