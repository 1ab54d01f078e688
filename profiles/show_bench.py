#!/usr/bin/env python
"""One summary line per bench JSON file: usage show_bench.py file..."""
import json, sys
for f in sys.argv[1:]:
    try:
        j = json.loads([l for l in open(f).read().splitlines() if l.startswith('{')][-1])
    except Exception as e:
        print(f, 'ERR', e); continue
    bd = {k: round(v, 2) for k, v in (j.get('breakdown_ms') or {}).items() if isinstance(v, float)}
    print(f.split('/')[-1], 'N', j['n_gpus'], 'value', round(j['value']), 'ms', round(j['ms_per_step'], 3), 'e2e', round(j['e2e']['value']),
          'launches', j.get('gpu_launches'), (j.get('message_set') or {}).get('count'), (j.get('message_set') or {}).get('hash'), j.get('sharded'), bd,
          'frac', round(j['roofline']['frac'], 4), 'clk', (j.get('clocks') or {}).get('sm_mhz'))
