#!/bin/bash
# Short kernel-only bench line (no e2e, no CPU baseline): step, phase breakdown, roofline fraction, cut check.
timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline "$@" 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']
print(round(j['ms_per_step'],3), j['breakdown_ms_rank0'], 'frac', round(r['frac'],4), 'cut', r['cut_check_worst'], 'fallback', r['sequential_fallback'])"
