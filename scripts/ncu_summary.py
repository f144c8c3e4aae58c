"""Summarise an .ncu-rep (ncu --set full) per kernel: python scripts/ncu_summary.py gpurun_out/prof.ncu-rep"""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
idx = {h: i for i, h in enumerate(hdr)}
want = ['gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'smsp__inst_executed.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_bytes.sum', 'lts__t_sectors.sum', 'lts__t_requests.sum', 'derived__lts__lts2xbar_bytes.sum.per_second',
        'dram__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct',
        'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'smsp__sass_average_branch_targets_threads_uniform.pct']
for r in data:
    print('----', r[idx['Kernel Name']].strip(), 'id', r[idx['ID']] if 'ID' in idx else '')
    for w in want:
        if w in idx:
            print('  %-70s %s %s' % (w, r[idx[w]], units[idx[w]]))
    st = [(float(r[i]), h) for h, i in idx.items()
          if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('_per_issue_active.ratio')]
    print('  stalls/issue: ' + ', '.join('%s %.2f' % (h.split('stalled_')[1].replace('_per_issue_active.ratio', ''), v)
                                          for v, h in sorted(st, reverse=True)[:6]))
