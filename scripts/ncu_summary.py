"""Summarise gpurun_out/launches.csv and gpurun_out/prof_composite.ncu-rep into profiles/*.md.
Usage: python scripts/ncu_summary.py <tag>"""
import collections, csv, re, subprocess, sys
tag = sys.argv[1]
lines = open('gpurun_out/launches.csv').read().splitlines()
start = [i for i, l in enumerate(lines) if l.startswith('"ID"')][0]
rows = list(csv.DictReader(lines[start:]))
acc = collections.OrderedDict()
for r in rows:
    n = re.sub(r'\(.*', '', r['Kernel Name']).replace('void ', '')[-60:]
    a = acc.setdefault(n, [0, 0.0]); a[0] += 1; a[1] += float(r['Metric Value'])
tot = sum(v[1] for v in acc.values())
with open(f'profiles/r01_ncu_launches_6M_{tag}.md', 'w') as f:
    f.write(f'# ncu launch list, round 1, kernels "{tag}"\n\n')
    f.write('Command: `ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv python bench.py --P 6000000 --views 2 --steps 1 --warmup 1 --no-cpu-baseline`\n')
    f.write('(first 600 launches: scene upload, warm-up step and part of the timed step; per-launch times are cold-cache and serialised — compare shares, not absolutes).\n\n')
    f.write('| kernel | launches | total ms | share |\n|---|---|---|---|\n')
    for n, (c, t) in sorted(acc.items(), key=lambda kv: -kv[1][1])[:25]:
        f.write(f'| `{n}` | {c} | {t/1e6:.3f} | {100*t/tot:.1f}% |\n')
raw = subprocess.run(['ncu', '-i', 'gpurun_out/prof_composite.ncu-rep', '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]; idx = {h: i for i, h in enumerate(hdr)}
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__cycles_active.avg', 'sm__cycles_elapsed.max',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'smsp__inst_executed.sum', 'launch__registers_per_thread',
        'launch__waves_per_multiprocessor', 'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'lts__t_sector_hit_rate.pct', 'sm__inst_executed_pipe_xu.sum',
        'smsp__pcsamp_warps_issue_stalled_barrier', 'smsp__pcsamp_warps_issue_stalled_wait', 'smsp__pcsamp_warps_issue_stalled_long_scoreboard',
        'smsp__pcsamp_warps_issue_stalled_short_scoreboard', 'smsp__pcsamp_warps_issue_stalled_math_pipe_throttle',
        'smsp__pcsamp_warps_issue_stalled_not_selected', 'smsp__pcsamp_warps_issue_stalled_selected', 'smsp__pcsamp_warps_issue_stalled_branch_resolving']
with open(f'profiles/r01_ncu_full_composite_{tag}.md', 'w') as f:
    f.write(f'# ncu --set full, composite kernels "{tag}" (6 M Gaussians, 1408x376, view 0)\n\n')
    f.write('Command: `ncu --set full --clock-control none --import-source on -k regex:composite_(fwd|bwd)_kernel -s 2 -c 2 python bench.py --P 6000000 --views 2 --steps 1 --warmup 1 --no-cpu-baseline`\n\n')
    names = [r[idx['Kernel Name']][-45:] for r in rows[2:]]
    f.write('| metric | unit | ' + ' | '.join(names) + ' |\n|---|---|' + '---|' * len(names) + '\n')
    for w in want:
        if w in idx:
            f.write(f'| {w} | {units[idx[w]]} | ' + ' | '.join(r[idx[w]] for r in rows[2:]) + ' |\n')
# DRAM traffic per launch (read + write) of the captured kernels, for bench.py's roofline.traffic
import json, os
tr = {}
scale = {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
for r in rows[2:]:
    name = r[idx['Kernel Name']]
    key = 'composite_fwd' if 'composite_fwd' in name else 'composite_bwd' if 'composite_bwd' in name else None
    if key and key not in tr:
        tr[key] = (float(r[idx['dram__bytes_read.sum']]) * scale[units[idx['dram__bytes_read.sum']]]
                   + float(r[idx['dram__bytes_write.sum']]) * scale[units[idx['dram__bytes_write.sum']]])
old = {}
if os.path.exists('profiles/ncu_traffic.json'):
    old = json.load(open('profiles/ncu_traffic.json')).get('bytes_per_launch', {})
old.update(tr)
json.dump({'bytes_per_launch': old,
           'source': f'ncu --set full --clock-control none, dram__bytes_read.sum + dram__bytes_write.sum, one launch (view 0 of the 6 M scene), capture "{tag}"'},
          open('profiles/ncu_traffic.json', 'w'), indent=1)
print(open(f'profiles/r01_ncu_launches_6M_{tag}.md').read()[:1800])
print(open(f'profiles/r01_ncu_full_composite_{tag}.md').read())
