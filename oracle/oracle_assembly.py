"""CPU ORACLE for the consensus assembly of window calls -- TEST INFRASTRUCTURE ONLY.

Restates ``process_output`` (nanotensor/dataset.py:421-432): ``index2base`` per window (nanotensor/trim_signal.py:116-122),
``simple_assembly`` (chiron/chiron/utils/easy_assembler.py:11-37), argmax, ``index2base``.
Only ``tests/`` and ``__graft_entry__.smoke()`` may import this; the product path never does.

PARITY PINNED: tests/golden/assembly.npz holds the outputs of the reference's own functions (executed verbatim by
tests/golden/make_golden_assembly.py); this restatement reproduces both the vote matrix and the consensus exactly
(tests/test_assembly.py).

The reference leans on a standard-library algorithm, ``difflib.SequenceMatcher`` (CPython 3.12 here): it takes
``max(get_matching_blocks(), key=size)``.  For a second sequence shorter than 200 elements (no "autojunk") that block is
the longest common substring found by the top-level ``find_longest_match``: every block of a sub-range is a candidate
of the full range, so none found later can be as long and start earlier.  Its tie-break is the scan order of the
dynamic programme -- end position in ``a`` ascending, then in ``b`` ascending, strict improvement only -- restated in
``longest_block`` below without difflib (checked against difflib on random strings in the tests).  With no common
element the only block is the sentinel ``(len(a), len(b), 0)``.

A second sequence of 200 or more elements switches difflib's "autojunk" on (``SequenceMatcher.__chain_b``): every
element occurring more than ``len(b) // 100 + 1`` times in ``b`` is POPULAR and leaves the match index -- with four bases
that is normally every base.  ``find_longest_match`` then finds the longest run of matches made of non-popular elements
only (same scan order), extends it at both ends over equal elements (popular ones included), ``get_matching_blocks``
recurses to the left and to the right of each block, sorts the blocks, merges adjacent ones, appends the sentinel, and the
reference takes the first block of maximal size.  ``longest_block_autojunk`` restates exactly that; it is pinned by
tests/golden/assembly_long.npz (the reference's functions run on T = 300 windows of 200-300 bases) and checked against
difflib itself on random strings.
"""
import numpy as np

AUTOJUNK_LEN = 200


def window_bases(row):
    """index2base as data: the entries of a label row that are not -1, in order (values 0..4 = A C G T E)."""
    row = np.asarray(row)
    return row[row != -1].astype(np.int8)


def longest_block(a, b):
    """(i, j, k): a[i:i+k] == b[j:j+k] with k maximal; among those the one ending first in a, then first in b;
    (len(a), len(b), 0) when nothing matches."""
    if len(b) >= AUTOJUNK_LEN:
        return longest_block_autojunk(a, b)
    best = (len(a), len(b), 0)
    prev = np.zeros(len(b) + 1, dtype=np.int32)
    for i in range(len(a)):
        cur = np.zeros(len(b) + 1, dtype=np.int32)
        eq = b == a[i]
        cur[1:][eq] = prev[:-1][eq] + 1
        k = int(cur.max()) if len(b) else 0
        if k > best[2]:
            j = int(np.argmax(cur)) - 1                    # first j reaching k in this row
            best = (i - k + 1, j - k + 1, k)
        prev = cur
    return best


def _find_longest(a, b, popular, alo, ahi, blo, bhi):
    """difflib.SequenceMatcher.find_longest_match with the popular elements out of the index and no junk."""
    besti, bestj, bestsize = alo, blo, 0
    j2len = {}
    for i in range(alo, ahi):
        new = {}
        if not popular[int(a[i])]:
            for j in range(blo, bhi):                      # b2j[a[i]] in ascending order, clipped to [blo, bhi)
                if b[j] == a[i]:
                    k = new[j] = j2len.get(j - 1, 0) + 1
                    if k > bestsize:
                        besti, bestj, bestsize = i - k + 1, j - k + 1, k
        j2len = new
    while besti > alo and bestj > blo and a[besti - 1] == b[bestj - 1]:
        besti, bestj, bestsize = besti - 1, bestj - 1, bestsize + 1
    while besti + bestsize < ahi and bestj + bestsize < bhi and a[besti + bestsize] == b[bestj + bestsize]:
        bestsize += 1
    return besti, bestj, bestsize


def longest_block_autojunk(a, b):
    """max(SequenceMatcher(None, a, b).get_matching_blocks(), key=size) for len(b) >= 200 (5 symbols)."""
    n = len(b)
    counts = np.bincount(np.asarray(b, dtype=np.int64), minlength=8)
    popular = counts > n // 100 + 1
    queue, blocks = [(0, len(a), 0, n)], []
    while queue:
        alo, ahi, blo, bhi = queue.pop()
        i, j, k = _find_longest(a, b, popular, alo, ahi, blo, bhi)
        if k:
            blocks.append((i, j, k))
            if alo < i and blo < j:
                queue.append((alo, i, blo, j))
            if i + k < ahi and j + k < bhi:
                queue.append((i + k, ahi, j + k, bhi))
    blocks.sort()
    merged, (i1, j1, k1) = [], (0, 0, 0)
    for i2, j2, k2 in blocks:
        if i1 + k1 == i2 and j1 + k1 == j2:
            k1 += k2
        else:
            if k1:
                merged.append((i1, j1, k1))
            i1, j1, k1 = i2, j2, k2
    if k1:
        merged.append((i1, j1, k1))
    merged.append((len(a), n, 0))
    best = merged[0]
    for blk in merged[1:]:
        if blk[2] > best[2]:
            best = blk
    return best


def simple_assembly(windows):
    """easy_assembler.py:11-37 on integer base arrays: returns the int vote matrix [5, length]."""
    votes = {}
    pos, length = 0, 0

    def add(start, seg):                                   # add_count, easy_assembler.py:31-37
        if start < 0:
            seg, start = seg[-start:], 0
        for i, base in enumerate(seg):
            votes[(int(base), start + i)] = votes.get((int(base), start + i), 0) + 1

    for idx, cur in enumerate(windows):
        if idx == 0:
            add(0, cur)
            continue
        i, j, _ = longest_block(windows[idx - 1], cur)
        disp = i - j
        add(pos + disp, cur)
        pos += disp
        length = max(length, pos + len(cur))               # never raised by the first window (easy_assembler.py:17-19)
    out = np.zeros((5, length), dtype=np.int32)
    for (base, p), n in votes.items():
        if p < length:
            out[base, p] = n
    return out


def assemble(labels, read_off):
    """process_output for many reads: labels [n_windows, T] int (-1 padded), read_off [n_reads + 1] window offsets.
    Returns (consensus ASCII uint8 concatenated, cons_off [n_reads + 1], votes [5, total])."""
    letters = np.frombuffer(b"ACGTE", dtype=np.uint8)
    cons, counts, off = [], [], [0]
    for a, b in zip(read_off[:-1], read_off[1:]):
        m = simple_assembly([window_bases(r) for r in labels[a:b]])
        counts.append(m)
        cons.append(letters[np.argmax(m, axis=0)] if m.shape[1] else np.zeros(0, dtype=np.uint8))
        off.append(off[-1] + m.shape[1])
    return (np.concatenate(cons) if cons else np.zeros(0, np.uint8), np.asarray(off, dtype=np.int64),
            np.concatenate(counts, axis=1) if counts else np.zeros((5, 0), np.int32))
