"""Summarise an `ncu --page raw --csv` export: one block per distinct kernel with the metrics
that decide the roofline discussion.  Usage: ncu_summary.py raw.csv [out.csv]"""
import csv
import sys

WANT = ['gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_sectors_srcunit_tex_op_red.sum', 'lts__t_sectors_srcunit_tex_op_red.avg.pct_of_peak_sustained_elapsed',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'dram__bytes_read.sum', 'dram__bytes_write.sum']
STALL = 'smsp__average_warps_issue_stalled_'


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    cols = [i for i, h in enumerate(hdr) if h in WANT or (h.startswith(STALL) and h.endswith('_per_issue_active.ratio'))]
    seen = {}
    for r in rows[2:]:
        name = r[hdr.index('Kernel Name')].split('(')[0]
        seen.setdefault(name, []).append(r)
    out = []
    for name, rs in seen.items():
        r = rs[-1]
        print('----', name, f'({len(rs)} launches captured; last shown)')
        for i in cols:
            try:
                v = float(r[i].replace(',', ''))
            except ValueError:
                continue
            if hdr[i].startswith(STALL) and v < 0.3:
                continue
            print(f'   {hdr[i]:100s} {r[i]:>16s} {units[i]}')
            out.append((name, hdr[i], r[i], units[i]))
    if len(sys.argv) > 2:
        with open(sys.argv[2], 'w', newline='') as f:
            csv.writer(f).writerows([('kernel', 'metric', 'value', 'unit')] + out)


if __name__ == '__main__':
    main()
