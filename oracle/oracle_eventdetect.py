"""CPU ORACLE for MinKNOW-style event detection -- TEST INFRASTRUCTURE ONLY.

Restates ``minknow_event_detect`` (nanonet/nanonet/eventdetection/filters.py:299-322): ``compute_sum_sumsq``
(filters.c:57-72), ``compute_tstat`` (filters.c:130-195), ``short_long_peak_detector`` (filters.c:235-300 with the
``Detector`` initial state of filters.py:116-143) and ``_construct_events`` (filters.py:146-204) -- SURVEY.md section 8(f)
rank 4, second half.  Only ``tests/`` and ``__graft_entry__.smoke()`` may import this.

PARITY PINNED: the reference's C file compiles unmodified (oracle/build_ref.sh -> oracle/_ref/libnanonetfilters.so) and
tests/golden/event_detect.npz holds what the reference's own Python + C produce; this restatement equals them bit for bit
(events and peaks), and where oracle/_ref is present the tests also run it live on fresh inputs.

Arithmetic notes that decide the last bit: the cumulative sums are SEQUENTIAL float64 additions (np.cumsum accumulates
in order, like the C loop); nothing is fused (stock gcc -O2 on x86-64 has no FMA to contract into); the event variance
uses Python's ``ev_mean ** 2`` = libm ``pow`` (filters.py:190) -- kept as is here, whereas the CUDA kernel multiplies,
which can differ in the last bit of ``stdv`` when pow is not correctly rounded (tolerance stated in the tests).
"""
import numpy as np

DEF_PEAK_POS, DEF_PEAK_VAL = -1, 1e100


def compute_sum_sumsq(data):
    data = np.asarray(data, dtype=np.float64)
    return np.cumsum(data), np.cumsum(data * data)


def compute_tstat(sums, sumsqs, w):
    """filters.c:130-195, pooled = false."""
    n = len(sums)
    t = np.zeros(n)
    if n < 2 * w or w < 2:
        return t
    i = np.arange(w, n - w + 1)
    sum1, sumsq1 = sums[i - 1].copy(), sumsqs[i - 1].copy()
    later = i > w
    sum1[later] -= sums[i[later] - w - 1]
    sumsq1[later] -= sumsqs[i[later] - w - 1]
    sum2 = sums[i + w - 1] - sums[i - 1]
    sumsq2 = sumsqs[i + w - 1] - sumsqs[i - 1]
    mean1, mean2 = sum1 / w, sum2 / w
    var1 = np.maximum(sumsq1 / w - mean1 * mean1, 1e-100)
    var2 = np.maximum(sumsq2 / w - mean2 * mean2, 1e-100)
    t[i] = np.abs((mean2 - mean1) / np.sqrt(var1 / w + var2 / w))
    return t


class _Det(object):
    def __init__(self, signal, threshold, window_length):
        self.signal, self.threshold, self.window_length = signal, threshold, window_length
        self.masked_to, self.peak_pos, self.peak_value, self.valid_peak = 0, DEF_PEAK_POS, DEF_PEAK_VAL, False


def short_long_peak_detector(t_stats, thresholds, window_lengths, peak_height):
    """filters.c:235-300; returns the emitted peak positions in emission order, zeros dropped (filters.py:240-241)."""
    dets = [_Det(*x) for x in zip(t_stats, thresholds, window_lengths)]
    if dets[0].window_length > dets[1].window_length:
        dets = dets[::-1]
    short, long_ = dets
    peaks = []
    for i in range(len(short.signal)):
        for d in dets:
            if d.masked_to >= i:
                continue
            cur = d.signal[i]
            if d.peak_pos == DEF_PEAK_POS:
                if cur < d.peak_value:
                    d.peak_value = cur
                elif cur - d.peak_value > peak_height:
                    d.peak_value, d.peak_pos = cur, i
            else:
                if cur > d.peak_value:
                    d.peak_value, d.peak_pos = cur, i
                if d is short and d.peak_value > d.threshold:
                    long_.masked_to = d.peak_pos + d.window_length
                    long_.peak_pos, long_.peak_value, long_.valid_peak = DEF_PEAK_POS, DEF_PEAK_VAL, False
                if d.peak_value - cur > peak_height and d.peak_value > d.threshold:
                    d.valid_peak = True
                if d.valid_peak and (i - d.peak_pos) > d.window_length // 2:
                    peaks.append(d.peak_pos)
                    d.peak_pos, d.peak_value, d.valid_peak = DEF_PEAK_POS, cur, False
    return np.asarray([p for p in peaks if p != 0], dtype=np.int64)


def construct_events(sums, sumsqs, peaks, sample_rate):
    """filters.py:146-204: [start, length, mean, stdv] per event, float64 [n_events, 4]."""
    n = len(sums)
    edges = [int(p) for p in peaks if 0 < p < n] + [n]
    ev = np.empty((len(edges), 4))
    s, sm, smsq = 0, 0, 0
    for k, e in enumerate(edges):
        ln = e - s
        mean = float(sums[e - 1] - sm) / ln
        var = max(0.0, (sumsqs[e - 1] - smsq) / ln - (mean ** 2))
        ev[k] = (s, ln, mean, np.sqrt(var))
        s, sm, smsq = e, sums[e - 1], sumsqs[e - 1]
    ev[:, 0] /= sample_rate
    ev[:, 1] /= sample_rate
    return ev


def minknow_event_detect(raw, sample_rate, window_lengths=(16, 40), thresholds=(8.0, 4.0), peak_height=1.0):
    sums, sumsqs = compute_sum_sumsq(raw)
    tstats = [compute_tstat(sums, sumsqs, int(w)) for w in window_lengths]
    peaks = short_long_peak_detector(tstats, list(thresholds), [int(w) for w in window_lengths], peak_height)
    return construct_events(sums, sumsqs, peaks, sample_rate), peaks
