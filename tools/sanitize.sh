#!/usr/bin/env bash
# compute-sanitizer evidence (memcheck / racecheck / synccheck on the mbarrier + TMEM protocols).
# Runs tools/sanitize_target.py (small forwards of every tensor-core path + the widened rows, each checked against its
# golden) under the three tools and writes gpurun_out/sanitizer_<tool>_<part>.txt; copy the summaries to profiles/.
#   gpurun --timeout 1500 -- 'bash tools/sanitize.sh'
set -uo pipefail
cd "$(dirname "${BASH_SOURCE[0]}")/.."
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
for tool in memcheck synccheck racecheck; do
  for part in forward stress rows; do
    out="gpurun_out/sanitizer_${tool}_${part}.txt"
    timeout "${SAN_TIMEOUT:-420}" "$CS" --tool "$tool" --print-limit 20 --error-exitcode 9 \
        python tools/sanitize_target.py "$part" > "$out" 2>&1
    rc=$?
    echo "exit code $rc" >> "$out"
    echo "$tool $part: rc $rc; $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' "$out" | tail -1)"
  done
done
