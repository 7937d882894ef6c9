"""Print the numbers of a bench.py JSON line that one looks at while tuning."""
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
for k in ('value', 'ms_per_step', 'lm_iterations_ran', 'accept_reject', 'linear_solver_iterations', 'final_cost', 'steps_run', 'parity_vs_1gpu'):
    if k in d: print(k, d[k])
for k in ('e2e', 'e2e_pageable', 'cpu_baseline'):
    if k in d: print(k, {a: b for a, b in d[k].items() if a not in ('unit', 'host_memory')})
if 'roofline' in d: print('roofline', {a: b for a, b in d['roofline'].items() if a != 'note'})
def table(t):
    for k, v in t.items():
        print('   %-14s' % k, {a: (round(b, 4) if isinstance(b, float) else b) for a, b in v.items()})
if 'stages' in d: table(d['stages'])
for n, e in d.get('extra_workloads', {}).items():
    print(n, {k: e[k] for k in e if k not in ('stages', 'config')})
    table(e['stages'])
