#!/usr/bin/env bash
# A/B timing of ablation switches on one box: stage_ms_per_step of bench.py at 65 536 windows per setting.
cd "$(dirname "$0")/.."
run() { echo "== $*"; env "$@" timeout 200 python bench.py --windows 65536 --steps 3 --warmup 2 --no-cpu --no-e2e 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']/1e6,2),'M/s', {k:round(v,2) for k,v in d['roofline']['stage_ms_per_step'].items()}, d['clocks']['sm_mhz'], 'launches', d['gpu_launches'], 'proj_gemm_us', round(d['roofline']['us_per_launch'],1))
    elif l: print(l[:300])
"; }
for cfg in "$@"; do run $cfg; done
