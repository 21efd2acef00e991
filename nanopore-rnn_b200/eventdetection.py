"""Host mirror of nanonet/nanonet/eventdetection/filters.py for the event-detection path:
``minknow_event_detect(raw_data, sample_rate, window_lengths, thresholds, peak_height)`` with the reference's name,
argument meaning and return value (structured array with float fields start, length, mean, stdv), computed by
``ntb_event_detect_host`` in libnanotensor_b200.so.  No CPU fallback; without the library / a B200 the call raises.
``minknow_event_detect_batch`` runs many reads in one call (one warp per read in the sequential stages: the GPU needs
thousands of reads in flight to be busy)."""
import ctypes as C

import numpy as np

from . import _lib

EVENT_DTYPE = [("start", float), ("length", float), ("mean", float), ("stdv", float)]


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def minknow_event_detect_batch(raws, sample_rate, window_lengths=[16, 40], thresholds=[8.0, 4.0], peak_height=1.0,
                               stream=None):
    """``raws``: sequence of 1-D sample arrays.  Returns a list of structured event arrays, one per read."""
    if sample_rate is None:
        raise NotImplementedError("minknow_event_detect: sample_rate=None (integer start/length) is not built")
    if len(window_lengths) != 2 or len(thresholds) != 2:
        raise ValueError("minknow_event_detect: two window lengths and two thresholds")
    raws = [np.ascontiguousarray(r, dtype=np.float64) for r in raws]
    if any(len(r) == 0 for r in raws):
        raise ValueError("minknow_event_detect: empty read")
    read_off = np.zeros(len(raws) + 1, dtype=np.int64)
    read_off[1:] = np.cumsum([len(r) for r in raws])
    total = int(read_off[-1])
    raw = np.concatenate(raws) if raws else np.zeros(0)
    wl = np.asarray(window_lengths, dtype=np.int32)
    th = np.asarray(thresholds, dtype=np.float64)
    events = np.empty((max(total, 1), 4), dtype=np.float64)          # an event holds at least one sample
    ev_off = np.zeros(len(raws) + 1, dtype=np.int64)
    lib = _lib.load()
    _lib.check(lib.ntb_event_detect_host(_ptr(raw), _ptr(read_off), len(raws), total, float(sample_rate), _ptr(wl),
                                         _ptr(th), float(peak_height), total, _ptr(events), _ptr(ev_off),
                                         C.c_void_p(stream or 0)))
    out = []
    for a, b in zip(ev_off[:-1], ev_off[1:]):
        ev = np.empty(b - a, dtype=EVENT_DTYPE)
        for k, name in enumerate(("start", "length", "mean", "stdv")):
            ev[name] = events[a:b, k]
        out.append(ev)
    return out


def minknow_event_detect(raw_data, sample_rate, window_lengths=[16, 40], thresholds=[8.0, 4.0], peak_height=1.0,
                         get_peaks=False):
    """filters.py:299-322."""
    if get_peaks:
        raise NotImplementedError("minknow_event_detect: get_peaks=True is outside the accelerated path")
    return minknow_event_detect_batch([raw_data], sample_rate, window_lengths, thresholds, peak_height)[0]
