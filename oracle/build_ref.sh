#!/usr/bin/env bash
# Compiles the REFERENCE'S OWN event-detection C source where it lies (nothing is copied into the repo):
#   /root/reference/nanonet/nanonet/eventdetection/filters.c  ->  oracle/_ref/libnanonetfilters.so
# (git-ignored; travels to the GPU box with the snapshot).  Plain gcc on the one file -- not the reference's setup.py.
# No -march / -ffast-math: the float64 arithmetic must stay un-fused, as in a stock build of the package.
# A no-op where /root/reference does not exist (the GPU box uses the prebuilt file).
set -euo pipefail
here="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
src=/root/reference/nanonet/nanonet/eventdetection/filters.c
[ -f "$src" ] || { echo "build_ref: $src not present, keeping any prebuilt oracle/_ref"; exit 0; }
mkdir -p "$here/_ref"
inc="$(python3 -c 'import sysconfig; print(sysconfig.get_paths()["include"])')"
gcc -O2 -w -shared -fPIC -I"$inc" -I/root/reference/nanonet/nanonet/include "$src" -o "$here/_ref/libnanonetfilters.so" -lm
echo "built $here/_ref/libnanonetfilters.so"
