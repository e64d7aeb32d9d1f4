"""Print the handful of ncu --set full metrics used in profiles/ from a .ncu-rep (run here, no GPU needed)."""
import csv, subprocess, sys, io
KEYS = ['gpu__time_duration.sum', 'launch__registers_per_thread', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_tensor.sum',
        'sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sectors_op_red.sum', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'smsp__thread_inst_executed_per_inst_executed.ratio']
def main(path, extra=()):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[0]
    for r in rows[2:]:
        print('==', r[hdr.index('Kernel Name')][:60])
        for i, h in enumerate(hdr):
            if h in KEYS or h.startswith('smsp__average_warps_issue_stalled') or any(e in h for e in extra):
                print('  %-90s %s %s' % (h, r[i], rows[1][i]))
if __name__ == '__main__':
    main(sys.argv[1], sys.argv[2:])
