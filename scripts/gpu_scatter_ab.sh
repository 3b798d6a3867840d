#!/bin/bash
set -u
for variant in "$@"; do
  env $variant python bench.py --cpu-fields 2 --e2e-steps 1 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$variant', d['config']['path'], round(d['value']), d['ms_per_step'], d['roofline']['frac'], d['parity_vs_cpu_sample'])"
done
