#!/bin/bash
# event detection: ms per call for every form of the sequential stages (NTB_ED_LANES = reads per warp, 0 = one warp per read)
for shape in "1024 50000" "4096 50000" "8192 50000" "16384 25000" "32768 12500"; do
  set -- $shape
  for r in 0 4 8 16 32; do
    NTB_ED_LANES=$r python tools/bench_eventdetect.py --reads $1 --samples $2 --iters 3 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('reads $1 x $2  R=$r  %s' % json.dumps({k: d[k] for k in d if k in ('value', 'ms_per_call', 'identical', 'same', 'unit')}))"
  done
done
