#!/usr/bin/env bash
# SASS evidence: per-kernel instruction histogram of the tensor-core / TMA / TMEM mnemonics in the built library
# (UTCHMMA = tcgen05.mma kind::f16, UTCBAR = tcgen05.commit, LDTM / STTM = tcgen05.ld / st, UTMALDG / UTMASTG = TMA
# tensor load / store, SYNCS = mbarrier ops).  Runs without a GPU (cuobjdump).  Output: profiles/r02_sass_histogram.txt
set -euo pipefail
cd "$(dirname "${BASH_SOURCE[0]}")/.."
so=nanopore-rnn_b200/libnanotensor_b200.so
out="${1:-profiles/r02_sass_histogram.txt}"
/usr/local/cuda/bin/cuobjdump -sass "$so" | python3 -c '
import re, sys, collections
pat = re.compile(r"\b(UTCHMMA(?:\.2CTA)?|UTCQMMA|UTCBAR(?:\.2CTA)?(?:\.MULTICAST)?|LDTM|STTM|UTMALDG|UTMASTG|UTMAPF|UTCATOMSWS|SYNCS|MUFU\.EX2|MUFU\.RCP|HMMA|FFMA|DADD|DMUL|DFMA)\b")
fn, hist, total = None, collections.OrderedDict(), {}
for line in sys.stdin:
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = m.group(1); hist[fn] = collections.Counter(); total[fn] = 0; continue
    if fn and re.match(r"\s+/\*[0-9a-f]{4,}\*/", line):
        total[fn] += 1
        for k in pat.findall(line):
            hist[fn][k] += 1
import subprocess
for fn, h in hist.items():
    name = subprocess.run(["/usr/local/cuda/bin/cu++filt", fn], capture_output=True, text=True).stdout.strip()
    if name.endswith(")"):                     # drop the parameter list (matching parenthesis of the last one)
        depth, i = 0, len(name) - 1
        while i >= 0:
            depth += name[i] == ")"
            depth -= name[i] == "("
            if depth == 0:
                break
            i -= 1
        name = name[:i]
    name = name.replace("ntb::", "").replace("(anonymous namespace)::", "").replace("<unnamed>::", "").replace("void ", "")
    keys = ", ".join("%s %d" % kv for kv in sorted(h.items()))
    print("%-62s %6d SASS instructions | %s" % (name[:62], total[fn], keys or "-"))
' > "$out"
echo "wrote $out"; grep -E "UTCHMMA|UTMALDG" "$out" | head -20
