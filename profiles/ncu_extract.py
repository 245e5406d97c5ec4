"""Print the ncu metrics we track, one line per kernel, from a `--page raw --csv` export.
usage: python profiles/ncu_extract.py raw.csv [--all]   (default: first launch of each kernel)"""
import csv
import sys

WANT = {'gpu__time_duration.sum': 'time', 'launch__registers_per_thread': 'regs', 'launch__shared_mem_per_block_dynamic': 'dsmem',
        'launch__grid_size': 'grid', 'sm__warps_active.avg.pct_of_peak_sustained_active': 'warps_act%', 'smsp__inst_executed.sum': 'inst',
        'smsp__thread_inst_executed_per_inst_executed.ratio': 'lanes', 'smsp__issue_active.avg.pct': 'issue%',
        'sm__inst_executed.avg.per_cycle_active': 'ipc', 'dram__bytes_read.sum': 'rd', 'dram__bytes_write.sum': 'wr',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed': 'dram%',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum': 'conf', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum': 'wavef',
        'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active': 'alu%', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active': 'fma%',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active': 'xu%', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active': 'lsu%',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio': 'st_long',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio': 'st_short',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio': 'st_bar',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio': 'st_wait',
        'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio': 'st_mio',
        'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio': 'st_notsel',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio': 'st_math',
        'smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio': 'st_br',
        'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio': 'st_noinst',
        'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio': 'st_lg',
        'smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio': 'st_disp'}


def rows_of(path):
    rows = list(csv.reader(open(path)))
    idx = {h: i for i, h in enumerate(rows[0])}
    units = rows[1]
    for r in rows[2:]:
        name = r[idx['Kernel Name']].split('(')[0].replace('b2::', '').replace('void ', '')
        d = {'kernel': name}
        for k, v in WANT.items():
            if k in idx:
                x = r[idx[k]].replace(',', '')
                try:
                    d[v] = float(x)
                except ValueError:
                    d[v] = x
                if v in ('time', 'rd', 'wr', 'dsmem'):
                    d[v + '_unit'] = units[idx[k]]
        yield d


def main():
    seen = {}
    for d in rows_of(sys.argv[1]):
        seen[d['kernel']] = seen.get(d['kernel'], 0) + 1
        if seen[d['kernel']] > 1 and '--all' not in sys.argv:
            continue
        out = [d['kernel'][:24].ljust(24)]
        for k, v in d.items():
            if k == 'kernel' or k.endswith('_unit'):
                continue
            u = d.get(k + '_unit', '')
            out.append('%s=%s%s' % (k, ('%.4g' % v) if isinstance(v, float) else v, u))
        print(' '.join(out))


if __name__ == '__main__':
    main()
