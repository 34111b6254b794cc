import json, sys
for f in sys.argv[1:]:
    try:
        d = json.load(open(f))
    except Exception as ex:
        print(f, "unreadable", ex); continue
    print(f, 'value', round(d['value'], 2), 'ms/step', round(d['ms_per_step'], 1), 'e2e', round(d['e2e']['value'], 2),
          'launches', d.get('gpu_launches'), d.get('clocks'))
    if 'kernels' in d:
        for k, v in d['kernels'].items():
            print('   %-15s %8.3f ms  %9.1f GB/s  frac %.4f' % (k, v['ms_per_view'], v['gbs'], v['frac']))
        print('   raw', {k: round(v, 3) for k, v in d['kernel_ms_per_view_raw'].items()})
        print('   stats', d['workload_stats'], 'view frac', round(d['view_roofline']['frac'], 4), 'ms/view', round(d['view_roofline']['ms_per_view'], 2))
        print('   cpu', d.get('cpu_baseline'))
