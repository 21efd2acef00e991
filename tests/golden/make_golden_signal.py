#!/usr/bin/env python
"""Generates tests/golden/signal_text.npz by RUNNING THE REFERENCE'S OWN read_signal (build container only).

SURVEY.md section 8(f) rank 2: ``read_signal`` (chiron/chiron/chiron_input.py:220-230) parses a ``.signal`` file with
``int(x) for x in line.split()``.  The function's text is cut out of the reference file and executed on small files
written here; ``normalize=False`` gives the parsed integers (the z-score that follows is ntb_pack_windows' part and has
its own tests).  The stored vectors are the file bytes and what the reference returned: this row is PINNED.
"""
import os
import re
import tempfile
import textwrap

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/chiron/chiron/chiron_input.py"


def reference_read_signal():
    src = open(REF).read()
    m = re.search(r"^(def read_signal\(.*?)(?=^def |\Z)", src, re.S | re.M)
    ns = {"np": np}
    exec(compile(textwrap.dedent(m.group(1)), REF, "exec"), ns)
    return ns["read_signal"]


def main():
    read_signal = reference_read_signal()
    rng = np.random.default_rng(3)
    a = rng.integers(200, 900, 30000)
    files = [
        " ".join(map(str, a)) + "\n",                                        # one long line (the shipped format)
        "\n".join(" ".join(map(str, a[i:i + 17])) for i in range(0, 5000, 17)),   # many lines, no trailing newline
        "  12\t-7 +5 007\r\n\r\n 32767 -32768   0\x0b1\x0c2 \n\n",            # every ASCII whitespace, signs, zeros
        "",                                                                   # empty file
        "5",                                                                  # one token, no whitespace at all
        " ".join(map(str, rng.integers(0, 10, 8190))) + " 123",              # tokens straddling the 4096-byte chunks
    ]
    text, off, samples, read_off = b"", [0], [], [0]
    with tempfile.TemporaryDirectory() as d:
        for k, t in enumerate(files):
            path = os.path.join(d, "r%d.signal" % k)
            with open(path, "w", newline="") as f:
                f.write(t)
            got = read_signal(path, normalize=False)                          # the reference's own function
            raw = open(path, "rb").read()
            text += raw
            off.append(len(text))
            samples.extend(got)
            read_off.append(len(samples))
    np.savez_compressed(os.path.join(HERE, "signal_text.npz"), text=np.frombuffer(text, dtype=np.uint8),
                        file_off=np.asarray(off, dtype=np.int64), samples=np.asarray(samples, dtype=np.int64),
                        read_off=np.asarray(read_off, dtype=np.int64))
    print("wrote signal_text.npz: %d files, %d bytes, %d samples" % (len(files), len(text), len(samples)))


if __name__ == "__main__":
    main()
