"""Host mirror of nanonet/nanonet/features.py for the event-feature front end:
``events_to_features(events, window=[-1, 0, 1], sloika_model=False)`` with the reference's name, argument meaning and
return value (float64 [n_events, 12]), computed by ``ntb_event_features_host`` in libnanotensor_b200.so -- bit-identical
to the reference.  There is no CPU fallback; without the library / a B200 the call raises.

``events_to_features_batch`` is the same for many reads in one launch (one CTA per read), which is the form a basecaller
run over a directory of reads should use (features.py:78-110 loops over files and calls events_to_features per read).
"""
import ctypes as C

import numpy as np

from . import _lib

WINDOW = (-1, 0, 1)
FIELDS = ("mean", "stdv", "length")


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def _field(events, name):
    return np.ascontiguousarray(events[name], dtype=np.float64)


def events_to_features_batch(reads, dtype=np.float64, stream=None):
    """``reads``: sequence of event tables (structured arrays / mappings with fields mean, stdv, length).
    Returns (features [total_events, 12] of ``dtype`` float64 or float32, read_off int64 [n_reads + 1])."""
    dtype = np.dtype(dtype)
    if dtype not in (np.dtype(np.float64), np.dtype(np.float32)):
        raise ValueError("events_to_features: dtype must be float64 or float32")
    lib = _lib.load()
    read_off = np.zeros(len(reads) + 1, dtype=np.int64)
    for r, ev in enumerate(reads):
        read_off[r + 1] = read_off[r] + len(ev["mean"])
    total = int(read_off[-1])
    cols = [np.concatenate([_field(ev, f) for ev in reads]) if reads else np.zeros(0) for f in FIELDS]
    out = np.empty((total, 4 * len(WINDOW)), dtype=dtype)
    _lib.check(lib.ntb_event_features_host(_ptr(cols[0]), _ptr(cols[1]), _ptr(cols[2]), _ptr(read_off), len(reads),
                                           total, 1 if dtype == np.float64 else 0, _ptr(out),
                                           C.c_void_p(stream or 0)))
    return out, read_off


def events_to_features(events, window=[-1, 0, 1], sloika_model=False):
    """features.py:61-75.  Only the path the event model uses is built: ``window`` [-1, 0, 1] and
    ``sloika_model=False``; anything else raises instead of silently computing something different."""
    if sloika_model:
        raise NotImplementedError("events_to_features: sloika_model=True is outside the accelerated path")
    if tuple(window) != WINDOW:
        raise NotImplementedError("events_to_features: only window=[-1, 0, 1] is built (got %r)" % (window,))
    return events_to_features_batch([events])[0]
