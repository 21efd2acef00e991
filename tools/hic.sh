#!/bin/bash
# end-to-end leg of bench.py with and without NVML polling inside it: per-step wall times (profiles/r02_e2e_nvml_hiccups.txt)
for i in 1 2 3; do for smp in ${MODES:-none thread step}; do
NTB_E2E_SAMPLER=$smp NTB_BENCH_DEBUG=1 python bench.py --steps 8 --warmup 1 --no-cpu --no-replica 2>&1 >/dev/null | grep "e2e step" | sed "s/device ms.*//; s/^/sampler=$smp /"
done; done
