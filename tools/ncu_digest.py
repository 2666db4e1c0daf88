#!/usr/bin/env python
"""Digest of an `ncu --page raw --csv` export: one row per captured launch."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
want = [('gpu__time_duration.sum', 'dur'), ('dram__bytes_read.sum', 'rd'), ('dram__bytes_write.sum', 'wr'),
        ('sm__warps_active.avg.pct_of_peak_sustained_active', 'occ%'), ('launch__registers_per_thread', 'regs'),
        ('sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm%'),
        ('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'dram%'),
        ('lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l2%'),
        ('smsp__thread_inst_executed_per_inst_executed.ratio', 'lanes'),
        ('sm__inst_executed.avg.per_cycle_elapsed', 'ipc'), ('smsp__inst_executed.sum', 'winst'),
        ('sm__icc_request_hit_rate.pct', 'icc%'),
        ('smsp__pcsamp_warps_issue_stalled_long_scoreboard', 'st_long'),
        ('smsp__pcsamp_warps_issue_stalled_wait', 'st_wait'),
        ('smsp__pcsamp_warps_issue_stalled_no_instructions', 'st_noinst'),
        ('smsp__pcsamp_warps_issue_stalled_short_scoreboard', 'st_short'),
        ('smsp__pcsamp_warps_issue_stalled_branch_resolving', 'st_branch'),
        ('smsp__pcsamp_warps_issue_stalled_selected', 'st_sel'),
        ('smsp__pcsamp_warps_issue_stalled_not_selected', 'st_notsel'),
        ('smsp__pcsamp_warps_issue_stalled_lg_throttle', 'st_lg'),
        ('smsp__pcsamp_warps_issue_stalled_barrier', 'st_bar'),
        ('smsp__pcsamp_warps_issue_stalled_math_pipe_throttle', 'st_math')]
ki = hdr.index('Kernel Name')
cols = [(hdr.index(m), n) for m, n in want if m in hdr]
print('| kernel | ' + ' | '.join(f'{n} [{units[i]}]' if units[i] else n for i, n in cols) + ' |')
print('|---|' + '---|' * len(cols))
for r in rows[2:]:
    vals = []
    for i, n in cols:
        try:
            vals.append(f"{float(r[i].replace(',', '')):.4g}")
        except ValueError:
            vals.append(r[i])
    print('| ' + r[ki].split('(')[0] + ' | ' + ' | '.join(vals) + ' |')

if len(sys.argv) > 3 and sys.argv[2] == '--traffic':
    # profiles/traffic.json: mean DRAM bytes per launch and kernel, scaled to a whole-shard launch
    import collections, json
    scale = float(sys.argv[4]) if len(sys.argv) > 4 else 1.0
    ir, iw = hdr.index('dram__bytes_read.sum'), hdr.index('dram__bytes_write.sum')
    mult = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
    acc = collections.defaultdict(list)
    for r in rows[2:]:
        name = r[ki].split('(')[0].replace('void ', '').split('<')[0].replace('hs::', '')
        acc[name].append(float(r[ir].replace(',', '')) * mult[units[ir]] + float(r[iw].replace(',', '')) * mult[units[iw]])
    # kernels not in this capture (k_agent_views has its own report) keep their previous entry
    try:
        out = json.load(open('profiles/traffic.json'))
    except (OSError, ValueError):
        out = {'kernels': {}}
    out['source'] = sys.argv[3]
    out['kernels'].update({k: scale * sum(v) / len(v) for k, v in acc.items()})
    json.dump(out, open('profiles/traffic.json', 'w'), indent=1)
