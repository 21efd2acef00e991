#!/usr/bin/env python
"""Generates tests/golden/assembly.npz by RUNNING THE REFERENCE'S OWN CODE (build container only; the GPU box reads the
committed output).

SURVEY.md section 8(f) rank 1: ``process_output`` (nanotensor/dataset.py:421-432) = ``SignalLabel.index2base``
(nanotensor/trim_signal.py:116-122) per window + ``simple_assembly`` (chiron/chiron/utils/easy_assembler.py:11-37) +
argmax + index2base.  easy_assembler.py is Python 2 (its ``main`` uses print statements), so the text of its two
functions (everything above ``def main``) is executed verbatim; index2base is cut out of trim_signal.py by its ``def``
block.  The vectors are therefore the reference's outputs: this row of the oracle is PINNED.

Inputs: label rows as the CTC readout produces them ([n_windows, T] int32, -1 padded): windows sliding over a random
base sequence with substitution / deletion errors, plus the edge cases -- a single-window read (the reference returns an
EMPTY consensus for it), empty windows, unrelated windows (no common block: displacement = len(prev) - len(cur)),
windows that drive the running position negative (clipped votes), a long read (> 1000 consensus positions: the padding
branch).
"""
import os
import re
import textwrap

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ASM = "/root/reference/chiron/chiron/utils/easy_assembler.py"
TRIM = "/root/reference/nanotensor/trim_signal.py"
T = 100


def reference_namespace():
    if not hasattr(np.lib, "pad"):          # NumPy 2 dropped the np.lib.pad alias the reference calls (it is np.pad)
        np.lib.pad = np.pad
    ns = {}
    src = open(ASM).read()
    exec(compile(src[:src.index("def main")], ASM, "exec"), ns)             # imports + simple_assembly + add_count
    trim = open(TRIM).read()
    m = re.search(r"^(    def index2base\(.*?)(?=^    def |\Z)", trim, re.S | re.M)
    exec(compile(textwrap.dedent(m.group(1)), TRIM, "exec"), ns)
    return ns


def sliding_read(rng, n_windows, step=8, width=26, err=0.06, T=T):
    truth = rng.integers(0, 4, step * n_windows + width + 8)
    rows = []
    for w in range(n_windows):
        seg = truth[w * step + rng.integers(0, 3): w * step + width + rng.integers(0, 5)].copy()
        flip = rng.random(len(seg)) < err
        seg[flip] = rng.integers(0, 5, int(flip.sum()))                     # substitutions, sometimes 'E' (4)
        seg = seg[rng.random(len(seg)) >= err / 2]                          # deletions
        rows.append(seg)
    return rows


def write_long(ns):
    """assembly_long.npz: T = 300 label rows with 200 and more called bases per window -- difflib's "autojunk" regime
    (len(b) >= 200: every element occurring more than len(b) // 100 + 1 times is dropped from the match index; with four
    bases that is every base, unless one is rare in the window).  Same reference functions, same layout as assembly.npz."""
    T2 = 300
    rng = np.random.default_rng(20171105)
    reads = []
    # (1) sliding windows of ~230 bases over one sequence: all four bases popular -> no anchors at all -> sentinel blocks
    reads.append(sliding_read(rng, 8, step=40, width=228, err=0.03, T=T2))
    # (2) the same with a RARE fifth symbol ('E' = 4) sprinkled in: the few E's are the only anchors, matches grow from them
    r = sliding_read(rng, 10, step=35, width=215, err=0.0, T=T2)
    truth_e = rng.integers(0, 4, 35 * 10 + 215 + 8)
    truth_e[rng.choice(len(truth_e), 9, replace=False)] = 4
    r = [truth_e[w * 35: w * 35 + 215 + int(rng.integers(0, 20))].copy() for w in range(10)]
    reads.append(r)
    # (3) long and short windows alternating (autojunk depends on the SECOND sequence's length only), a rare base in some
    r = []
    for w in range(12):
        n = int(rng.integers(200, 300)) if w % 2 else int(rng.integers(20, 199))
        seg = rng.integers(0, 3, n)                       # A, C, G popular
        pos = rng.choice(n, int(rng.integers(0, 4)), replace=False)
        seg[pos] = 3                                      # at most 3 T's: never popular in a long window
        r.append(seg)
    reads.append(r)
    # (4) exactly 199 / 200 / 201 bases around the threshold, related windows
    base = rng.integers(0, 4, 260)
    reads.append([base[:199].copy(), base[5:205].copy(), base[3:204].copy(), base[10:209].copy(), base[0:300].copy(), base[40:250].copy()])
    labels, lens, off, cons_txt, counts, cons_off = [], [], [0], [], [], [0]
    for rows in reads:
        block = np.full((len(rows), T2), -1, dtype=np.int32)
        for i, seg in enumerate(rows):
            block[i, :len(seg)] = seg
        all_reads = [ns["index2base"](row) for row in block]
        cons = ns["simple_assembly"](all_reads)
        txt = ns["index2base"](np.argmax(cons, axis=0)) if cons.shape[1] else ""
        labels.append(block)
        lens.extend(len(sg) for sg in rows)
        off.append(off[-1] + len(rows))
        cons_txt.append(txt)
        counts.append(cons.astype(np.int32))
        cons_off.append(cons_off[-1] + cons.shape[1])
    np.savez_compressed(os.path.join(HERE, "assembly_long.npz"), labels=np.concatenate(labels),
                        label_len=np.asarray(lens, dtype=np.int32), read_off=np.asarray(off, dtype=np.int64),
                        consensus=np.frombuffer("".join(cons_txt).encode(), dtype=np.uint8),
                        counts=np.concatenate(counts, axis=1), cons_off=np.asarray(cons_off, dtype=np.int64))
    print("wrote assembly_long.npz: %d reads, %d windows (%d with >= 200 bases), consensus lengths %s"
          % (len(reads), off[-1], sum(l >= 200 for l in lens), np.diff(cons_off)))


def main():
    ns = reference_namespace()
    rng = np.random.default_rng(20171104)
    reads = [sliding_read(rng, 60), sliding_read(rng, 1), sliding_read(rng, 2), sliding_read(rng, 150, step=9)]
    r = sliding_read(rng, 12)
    r[3] = r[3][:0]
    r[4] = r[4][:0]
    r[9] = r[9][:1]
    reads.append(r)                                                         # empty and one-base windows
    reads.append([rng.integers(0, 4, rng.integers(5, 30)) for _ in range(20)])   # unrelated windows
    back = sliding_read(rng, 30)[::-1]                                      # sliding backwards: position goes negative
    reads.append(back)
    reads.append([np.full(40, 2), np.full(12, 2), np.full(40, 2), np.arange(40) % 4, (np.arange(50) + 1) % 4])  # ties
    reads.append([rng.integers(0, 4, 99) for _ in range(3)] + [rng.integers(0, 4, 100)])   # full-width rows

    labels, lens, off = [], [], [0]
    cons_txt, counts, cons_off = [], [], [0]
    for rows in reads:
        block = np.full((len(rows), T), -1, dtype=np.int32)
        for i, seg in enumerate(rows):
            block[i, :len(seg)] = seg
        all_reads = [ns["index2base"](row) for row in block]               # dataset.py:427-428
        cons = ns["simple_assembly"](all_reads)                            # dataset.py:429
        txt = ns["index2base"](np.argmax(cons, axis=0)) if cons.shape[1] else ""   # dataset.py:430
        labels.append(block)
        lens.extend(len(s) for s in rows)
        off.append(off[-1] + len(rows))
        cons_txt.append(txt)
        counts.append(cons.astype(np.int32))
        cons_off.append(cons_off[-1] + cons.shape[1])
    write_long(ns)
    np.savez_compressed(os.path.join(HERE, "assembly.npz"), labels=np.concatenate(labels),
                        label_len=np.asarray(lens, dtype=np.int32), read_off=np.asarray(off, dtype=np.int64),
                        consensus=np.frombuffer("".join(cons_txt).encode(), dtype=np.uint8),
                        counts=np.concatenate(counts, axis=1), cons_off=np.asarray(cons_off, dtype=np.int64))
    print("wrote assembly.npz: %d reads, %d windows, consensus lengths %s" % (len(reads), off[-1], np.diff(cons_off)))


if __name__ == "__main__":
    main()
