"""Dev tool: print the interesting keys of bench.py JSON lines on stdin."""
import json, sys
for ln in sys.stdin:
    if ln.startswith('{'):
        d = json.loads(ln)
        print({k: (round(v, 1) if isinstance(v, float) else v) for k, v in d.items()
               if k in ('impl', 'value', 'ms_per_step', 'encode_mbins_s', 'decode_mbins_s', 'gpu_launches')})
        if 'roofline' in d: print('  kernel_ms', d['roofline']['kernel_ms'], 'frac', round(d['roofline']['frac'], 5))
        if 'issue_roofline' in d: print('  issue', d['issue_roofline'])
        if 'e2e' in d: print('  e2e', d['e2e'].get('value'), d['e2e'].get('ms_per_step'))
        if 'cpu_baseline' in d: print('  cpu', d['cpu_baseline']['value'], d['cpu_baseline'].get('single_thread_mbins_s'))
        if 'clocks' in d: print('  clocks', d['clocks'])
    else:
        print(ln.rstrip())
