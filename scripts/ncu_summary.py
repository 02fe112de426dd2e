"""Summarise an `ncu --set full` report (.ncu-rep) as one text block per captured launch.

    python scripts/ncu_summary.py gpurun_out/<tag>/prof_c4.ncu-rep > profiles/<tag>_prof_c4.txt

Reads the report with `ncu -i ... --page raw --csv` (works without a GPU).
"""
import csv
import io
import subprocess
import sys

WANT = [
    ('gpu__time_duration.sum', 'duration'),
    ('launch__grid_size', 'grid'),
    ('launch__block_size', 'block'),
    ('launch__registers_per_thread', 'regs/thread'),
    ('launch__shared_mem_per_block_dynamic', 'dyn smem/block'),
    ('launch__occupancy_limit_registers', 'occ limit regs (blocks)'),
    ('launch__occupancy_limit_shared_mem', 'occ limit smem (blocks)'),
    ('sm__warps_active.avg.pct_of_peak_sustained_active', 'warps active %'),
    ('dram__bytes_read.sum', 'dram read'),
    ('dram__bytes_write.sum', 'dram write'),
    ('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'dram throughput % of ncu peak'),
    ('lts__t_bytes.sum', 'L2 bytes'),
    ('lts__t_sector_hit_rate.pct', 'L2 hit %'),
    ('sm__throughput.avg.pct_of_peak_sustained_elapsed', 'SM throughput %'),
    ('smsp__issue_active.avg.pct_of_peak_sustained_active', 'issue slots active %'),
    ('sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active', 'fp64 pipe %'),
    ('sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'fma pipe %'),
    ('sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'xu (sfu) pipe %'),
    ('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'tensor pipe cycles active %'),
    ('sm__pipe_tensor_subpipe_cycles_active.avg.pct_of_peak_sustained_active', 'tensor subpipe active %'),
    ('sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active', 'uniform pipe %'),
    ('l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'smem wavefronts'),
    ('smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio', 'stall long scoreboard'),
    ('smsp__average_warp_latency_issue_stalled_short_scoreboard.ratio', 'stall short scoreboard'),
    ('smsp__average_warp_latency_issue_stalled_wait.ratio', 'stall wait'),
    ('smsp__average_warp_latency_issue_stalled_math_pipe_throttle.ratio', 'stall math pipe throttle'),
    ('smsp__average_warp_latency_issue_stalled_barrier.ratio', 'stall barrier'),
    ('smsp__average_warp_latency_issue_stalled_lg_throttle.ratio', 'stall lg throttle'),
    ('smsp__average_warp_latency_issue_stalled_not_selected.ratio', 'stall not selected'),
]


def main(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    ki = hdr.index('Kernel Name')
    cols = [(hdr.index(m), m, label) for m, label in WANT if m in hdr]
    tensor_like = [i for i, h in enumerate(hdr) if h in (
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed',
        'sm__ops_path_tensor_src_tf32_dst_fp32.avg.pct_of_peak_sustained_elapsed',
        'sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_active')]
    print('# %s' % path)
    for r in rows[2:]:
        print('kernel: %s' % r[ki][:150])
        for i, m, label in cols:
            print('  %-34s %14s %-8s (%s)' % (label, r[i], units[i], m))
        for i in tensor_like:
            if (i, hdr[i]) not in [(c[0], c[1]) for c in cols] and r[i] not in ('', '0', 'n/a'):
                print('  %-34s %14s %-8s (%s)' % ('tensor metric', r[i], units[i], hdr[i]))
        print()


if __name__ == '__main__':
    main(sys.argv[1])
