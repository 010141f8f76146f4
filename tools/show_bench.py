"""Readable summary of a bench.py JSON line:  python tools/show_bench.py file.json"""
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
def show(e, ind=0):
    pad = ' ' * ind
    for k in ('value', 'ms_per_step', 'windows', 'value_fp32', 'ms_per_step_fp32', 'gpu_launches', 'clocks', 'error', 'trace', 'single_gpu'):
        if k in e: print(pad, k, e[k])
    if 'config' in e: print(pad, 'path', e['config'].get('device_path'), '|', e['config'].get('l2'), '|', e['config'].get('workload'))
    for k in ('parity', 'parity_fp32', 'cpu_baseline'):
        if k in e: print(pad, k, {a: b for a, b in e[k].items() if a not in ('err_is', 'per_step')} if isinstance(e[k], dict) else e[k])
    if 'e2e' in e: print(pad, 'e2e', {k: v for k, v in e['e2e'].items() if k not in ('api', 'blocking_api')})
    if 'graph_build' in e: print(pad, 'graph_build', e['graph_build'])
    r = e.get('roofline')
    if r:
        print(pad, 'roofline kernel', r.get('kernel'), 'frac', r.get('frac'), 'fwd', {k: v for k, v in (r.get('forward') or {}).items() if k != 'note'})
        for k, v in (r.get('kernels') or {}).items(): print(pad, '  ', k, {a: (round(b, 4) if isinstance(b, float) else b) for a, b in v.items() if a != 'note'})
        for k in ('stages_ms_max_over_ranks', 'exchange', 'host_enqueue_ms_per_forward'):
            if k in r: print(pad, '  ', k, r[k])
show(d)
for n, e in d.get('workloads', {}).items():
    print('====', n); show(e, 2)
