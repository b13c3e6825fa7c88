"""ncu-rep -> short text summary per kernel launch (the metrics DESIGN.md / bench.py quote).
usage: python tools/ncu_summary.py <file.ncu-rep> [max launches per kernel name]"""
import csv
import subprocess
import sys

METRICS = [
    'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
    'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct',
    'l1tex__t_sector_hit_rate.pct', 'lts__t_sectors_op_red.sum', 'lts__t_sectors_op_atom.sum',
    'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__issue_active.avg.pct',
    'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
    'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
    'launch__grid_size', 'launch__block_size', 'smsp__inst_executed.sum',
    'smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio',
    'smsp__average_warp_latency_issue_stalled_lg_throttle.ratio',
    'smsp__average_warp_latency_issue_stalled_short_scoreboard.ratio',
    'smsp__average_warp_latency_issue_stalled_math_pipe_throttle.ratio',
    'smsp__average_warp_latency_issue_stalled_barrier.ratio',
    'smsp__average_warp_latency_issue_stalled_membar.ratio',
]


def main():
    rep = sys.argv[1]
    per_name = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    seen = {}
    for r in rows[2:]:
        name = r[hdr.index('Kernel Name')].split('(')[0].replace('<unnamed>::', '')
        seen[name] = seen.get(name, 0) + 1
        if seen[name] > per_name:
            continue
        print(f'== {name}  (launch {seen[name]})')
        for m in METRICS:
            if m in hdr:
                i = hdr.index(m)
                print(f'   {m:75s} {r[i]} {units[i]}')


if __name__ == '__main__':
    main()
