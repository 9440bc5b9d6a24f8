#!/usr/bin/env python
"""Summarise an .ncu-rep here (no GPU): key raw metrics per captured launch, plus the hottest source lines.
usage: python tools/ncu_summary.py gpurun_out/prof_X.ncu-rep [--source N]"""
import csv, io, subprocess, sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'dram__bytes_write.sum.per_second',
        'lts__t_sectors_srcunit_tex_op_write.sum', 'lts__t_sector_hit_rate.pct',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__grid_size',
        'launch__block_size', 'launch__shared_mem_per_block_dynamic', 'smsp__inst_executed.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_tensor.sum', 'smsp__average_warp_latency_per_inst_issued.ratio',
        'sm__cycles_elapsed.max', 'smsp__cycles_active.avg', 'l1tex__t_bytes_pipe_lsu_mem_global_op_st.sum',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'sm__pipe_tc_cycles_active.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed', 'sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active']


def raw(rep):
    out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    rows = raw(rep)
    hdr, units, data = rows[0], rows[1], rows[2:]
    ki = hdr.index('Kernel Name')
    print('kernels:', [r[ki][:60] for r in data])
    for i, h in enumerate(hdr):
        if h in KEYS or 'issue_stalled' in h and h.endswith('per_issue_active.ratio'):
            vals = [r[i] for r in data]
            try:
                if all(float(v.replace(',', '')) < 0.05 for v in vals) and 'stalled' in h:
                    continue
            except ValueError:
                pass
            print('%-90s %-12s %s' % (h, units[i], ' '.join(vals)))
    if '--source' in sys.argv:
        top = int(sys.argv[sys.argv.index('--source') + 1])
        out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'] , capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(out)))
        # find header
        for hi, r in enumerate(rows):
            if 'Source' in r and any('Sampl' in c for c in r):
                break
        else:
            print('no source page'); return
        h = rows[hi]
        si = h.index('Source')
        samp = [i for i, c in enumerate(h) if c.startswith('# Samples') or c == 'Warp Stall Sampling (All Samples)' or c == 'Samples']
        print('columns:', [c for c in h][:40])
        if not samp:
            return
        c0 = samp[0]
        body = [r for r in rows[hi + 1:] if len(r) > c0]
        def num(x):
            try: return float(x.replace(',', ''))
            except ValueError: return 0.0
        body.sort(key=lambda r: -num(r[c0]))
        tot = sum(num(r[c0]) for r in body) or 1
        for r in body[:top]:
            print('%6.2f%%  %s' % (100 * num(r[c0]) / tot, r[si][:150]))


if __name__ == '__main__':
    main()
