"""Consensus assembly of window calls on the GPU: the batched form of
``process_output`` (nanotensor/dataset.py:421-432) -- index2base, ``simple_assembly``
(chiron/chiron/utils/easy_assembler.py:11-37), argmax, index2base -- for all reads of a job in one call of
``ntb_assemble_reads_host``.  Results are identical to the reference's (integer work).  No CPU fallback: without the
library / a B200 the call raises.  ``dataset.FullSignalSequence.process_output`` keeps the reference's one-read
signature; ``RunTensorflow`` uses this module when it has the labels of many reads at once.
"""
import ctypes as C
import os

import numpy as np

from . import _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def assemble_reads(labels, read_off, want_counts=False, stream=None):
    """labels [n_windows, T] int (-1 padded rows as the CTC readout returns them), read_off [n_reads + 1] window offsets.
    Returns a list of consensus strings (and the list of [5, length] int32 vote matrices when ``want_counts``)."""
    labels = np.ascontiguousarray(labels, dtype=np.int32)
    if labels.ndim != 2:
        raise ValueError("assemble_reads: labels must be [n_windows, T]")
    read_off = np.ascontiguousarray(read_off, dtype=np.int64)
    n_windows, T = labels.shape
    n_reads = len(read_off) - 1
    capacity = int((labels != -1).sum())
    cons = np.empty(max(capacity, 1), dtype=np.uint8)
    cons_off = np.zeros(n_reads + 1, dtype=np.int64)
    counts = np.zeros(5 * max(capacity, 1), dtype=np.int32) if want_counts else None
    lib = _lib.load()
    _lib.check(lib.ntb_assemble_reads_host(_ptr(labels), _ptr(read_off), n_reads, n_windows, max(T, 1), capacity,
                                           _ptr(cons), _ptr(cons_off), _ptr(counts) if want_counts else None,
                                           C.c_void_p(stream or 0)))
    text = cons[:cons_off[-1]].tobytes().decode("ascii")
    out = [text[a:b] for a, b in zip(cons_off[:-1], cons_off[1:])]
    if not want_counts:
        return out
    total = int(cons_off[-1])
    m = counts[:5 * total].reshape(5, total)
    return out, [m[:, a:b] for a, b in zip(cons_off[:-1], cons_off[1:])]


def write_fastas(names, labels, read_off, output_dir):
    """dataset.py:423-432 for many reads: ``<name>.fasta`` with ``>name`` and the consensus; returns the paths."""
    paths = []
    for name, seq in zip(names, assemble_reads(labels, read_off)):
        path = os.path.join(output_dir, name + ".fasta")
        with open(path, "w+") as f:
            f.write(">{}\n{}\n".format(name, seq))
        paths.append(path)
    return paths
