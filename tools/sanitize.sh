#!/usr/bin/env bash
# compute-sanitizer evidence (memcheck / racecheck / synccheck on the mbarrier + TMEM protocols).
# Runs tools/sanitize_target.py (small forwards of every tensor-core path -- the fused BLSTM CTA-pair kernel included --
# + the widened rows, each checked against its golden) under ONE tool per invocation (the profiling recipe allows one
# compute-sanitizer tool per gpurun call) and writes gpurun_out/sanitizer_<tool>_<part>.txt; copy the summaries to profiles/.
#   gpurun --timeout 1500 -- 'bash tools/sanitize.sh memcheck'      (then synccheck, then racecheck, one call each)
set -uo pipefail
cd "$(dirname "${BASH_SOURCE[0]}")/.."
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
tool="${1:?usage: sanitize.sh memcheck|synccheck|racecheck}"
python tools/sanitize_target.py all > "gpurun_out/sanitizer_${tool}_plain.txt" 2>&1 || { echo "plain run failed"; tail -5 "gpurun_out/sanitizer_${tool}_plain.txt"; exit 1; }
for part in forward stress rows; do
  out="gpurun_out/sanitizer_${tool}_${part}.txt"
  timeout "${SAN_TIMEOUT:-420}" "$CS" --tool "$tool" --print-limit 20 --error-exitcode 9 \
      python tools/sanitize_target.py "$part" > "$out" 2>&1
  rc=$?
  echo "exit code $rc" >> "$out"
  echo "$tool $part: rc $rc; $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' "$out" | tail -1)"
done
