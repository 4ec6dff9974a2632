"""Summarise an ncu --set full report (.ncu-rep) into text: key raw metrics, stall reasons, opcode mix."""
import collections
import csv
import io
import re
import subprocess
import sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__throughput.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__grid_size',
        'launch__block_size', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
        'smsp__warps_eligible.avg.per_cycle_active', 'sass__inst_executed_register_spilling', 'inst_executed',
        'thread_inst_executed_true', 'smsp__sass_inst_executed_op_tma_ld.sum', 'sass__inst_executed_shared_loads',
        'sass__inst_executed_global_loads', 'sass__inst_executed_global_stores',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum']


def run(args):
    return subprocess.run(['ncu'] + args, capture_output=True, text=True).stdout


def main(rep, out=None):
    text = []
    rows = list(csv.reader(io.StringIO(run(['-i', rep, '--page', 'raw', '--csv']))))
    hdr, units, data = rows[0], rows[1], rows[2:]
    for d in data:
        text.append('== ' + d[hdr.index('Kernel Name')])
        for k in KEYS:
            if k in hdr:
                text.append('  %-70s %18s %s' % (k, d[hdr.index(k)], units[hdr.index(k)]))
        text.append('  stall reasons (warps per issue-active cycle):')
        for i, h in enumerate(hdr):
            if 'issue_stalled' in h and h.endswith('_per_issue_active.ratio') and 'not_issued' not in h:
                v = float(d[i])
                if v >= 0.05:
                    text.append('    %-40s %.2f' % (h.split('issue_stalled_')[1].replace('_per_issue_active.ratio', ''), v))
    src = list(csv.reader(io.StringIO(run(['-i', rep, '--page', 'source', '--csv']))))
    h2 = src[1]
    iS, iE = h2.index('Source'), h2.index('Instructions Executed')
    ops = collections.Counter()
    for r in src[2:]:
        if len(r) == len(h2) and r[iE].isdigit() and r[0].startswith('0x'):
            m = re.match(r'(@!?U?P\d+\s+)?([A-Z0-9_.]+)', r[iS].strip())
            ops[m.group(2).split('.')[0] if m else '?'] += int(r[iE])
    tot = sum(ops.values())
    text.append('== executed warp instructions by opcode (first kernel in the report), total %d' % tot)
    for op, c in ops.most_common(24):
        text.append('  %-10s %5.1f%%' % (op, 100.0 * c / tot))
    text = '\n'.join(text)
    print(text)
    if out:
        open(out, 'w').write(text + '\n')


if __name__ == '__main__':
    main(*sys.argv[1:])
