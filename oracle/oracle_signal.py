"""CPU ORACLE for the `.signal` text parse -- TEST INFRASTRUCTURE ONLY.

Restates the parse in ``read_signal`` (chiron/chiron/chiron_input.py:220-226): ``int(x) for x in line.split()`` over
the lines of the file.  Only ``tests/`` may import this.

PARITY PINNED: tests/golden/signal_text.npz holds what the reference's own read_signal returned for the file bytes
stored beside it (tests/golden/make_golden_signal.py); this restatement reproduces it exactly (tests/test_signal_parse.py).
The grammar is narrowed to what the kernel accepts -- [+-]?[0-9]{1,6} tokens between ASCII whitespace, values in int16
-- and anything else is an error here as there (Python's int() would also take underscores and non-ASCII digits).
"""
import re

import numpy as np

TOKEN = re.compile(rb"[+-]?[0-9]{1,6}\Z")
WS = b" \t\n\r\x0b\x0c\x1c\x1d\x1e\x1f"


def parse_signal(data):
    """bytes of one file -> int64 array of its samples."""
    out = []
    for tok in bytes(data).translate(bytes.maketrans(WS, b" " * len(WS))).split(b" "):
        if not tok:
            continue
        if not TOKEN.match(tok):
            raise ValueError("token %r is not [+-]?[0-9]{1,6}" % tok)
        v = int(tok)
        if not -32768 <= v <= 32767:
            raise ValueError("sample %d does not fit int16" % v)
        out.append(v)
    return np.asarray(out, dtype=np.int64)


def parse_signals(text, file_off):
    """Concatenated files (the C ABI's layout) -> (samples int64, read_off int64 [n_files + 1])."""
    parts = [parse_signal(text[a:b]) for a, b in zip(file_off[:-1], file_off[1:])]
    off = np.zeros(len(parts) + 1, dtype=np.int64)
    off[1:] = np.cumsum([len(p) for p in parts])
    return (np.concatenate(parts) if parts else np.zeros(0, np.int64)), off
