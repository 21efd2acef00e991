"""CPU ORACLE for the event-feature front end -- TEST INFRASTRUCTURE ONLY.

Restates ``nanonet/nanonet/features.py::events_to_features``: the 12 features per event
(mean, stdv, length, delta at event offsets -1, 0, +1) that the event (CrossEntropy) BLSTM model consumes.  Only
``tests/`` and ``__graft_entry__.smoke()`` may import this; the product path never does.

PARITY PINNED: tests/golden/event_features.npz holds the outputs of the reference's own functions (cut out of
/root/reference/nanonet/nanonet/features.py with ``ast`` and executed by tests/golden/make_golden_events.py; NumPy 2.3.x
in this image) and this restatement is checked against them BIT FOR BIT (tests/test_events.py).

The arithmetic that decides the last bit lives in a third-party dependency, NumPy: ``np.mean`` / ``np.std`` reduce with
PAIRWISE summation (numpy/_core/src/umath/loops_utils.h.src ``@TYPE@_pairwise_sum``; _methods.py ``_mean`` / ``_var``),
so that summation order is restated here explicitly (``pairwise_sum`` below, no call to np.sum / np.mean / np.std on the
path) -- it is the order the CUDA kernel follows, which is what makes the GPU result bit-identical as well.
"""
import numpy as np

PW_BLOCK = 128          # numpy's PW_BLOCKSIZE
ATOL = 1.0e-8           # np.allclose(x, 0.0): |x| <= atol + rtol * |0|


def pairwise_sum(a):
    """NumPy's float64 add-reduction order for one contiguous-or-strided 1-D run of n values:
    n < 8: left to right from 0.0; n <= 128: eight interleaved running sums r[j] += a[8k+j], folded as
    ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)), then the n%8 tail left to right; otherwise split at n//2 rounded down to a
    multiple of 8 and add the two halves."""
    n = len(a)
    if n < 8:
        res = np.float64(0.0)
        for v in a:
            res = res + v
        return res
    if n <= PW_BLOCK:
        r = [np.float64(a[j]) for j in range(8)]
        full = n - n % 8
        for i in range(8, full, 8):
            for j in range(8):
                r[j] = r[j] + a[i + j]
        res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]))
        for i in range(full, n):
            res = res + a[i]
        return res
    n2 = n // 2
    n2 -= n2 % 8
    return pairwise_sum(a[:n2]) + pairwise_sum(a[n2:])


_SUM = pairwise_sum      # tests switch this to np.add.reduce for reads too long for the pure-Python loop above


def _mean(x):
    """np.mean / ndarray.mean of a float64 vector (_methods.py _mean: umr_sum / count)."""
    return (np.float64(0.0) + _SUM(x)) / np.float64(len(x))      # the reduction starts from the identity 0.0


def _std(x):
    """np.std, ddof 0 (_methods.py _var / _std): mean, deviations, squares, pairwise sum, / n, sqrt."""
    d = x - _mean(x)
    d = d * d
    return np.sqrt((np.float64(0.0) + _SUM(d)) / np.float64(len(x)))


def scale_array(x, with_mean=True):
    """features.py:32-58 (with_std always true on this path): centre, re-centre when the residual mean is not
    ``allclose`` to 0, divide by the population std (1.0 when it is 0), re-centre again under the same test."""
    x = np.array(x, dtype=np.float64)
    if with_mean:
        x = x - _mean(x)
        m1 = _mean(x)
        if not abs(m1) <= ATOL:
            x = x - m1
    s = _std(x)
    if s == 0.0:
        s = np.float64(1.0)
    x = x / s
    if with_mean:
        m2 = _mean(x)
        if not abs(m2) <= ATOL:
            x = x - m2
    return x


def padded_offset(a, pos):
    """features.py:14-29: out[i] = a[i + pos], zero where that runs off either end."""
    out = np.zeros(len(a), dtype=np.float64)
    if pos == 0:
        out[:] = a
    elif pos > 0:
        out[:len(a) - pos] = a[pos:]
    else:
        out[-pos:] = a[:pos]
    return out


def events_to_features(mean, stdv, length, window=(-1, 0, 1)):
    """features.py:61-75 + SquiggleFeatureGenerator.__init__ (features.py:266-293, sloika_model=False): the three event
    fields standardised over the read; delta = first difference of the STANDARDISED mean (0 in front), divided by its
    population std without centring; columns ordered pos-major: mean, stdv, length, delta for each pos in ``window``.
    Returns float64 [n, 4 * len(window)]."""
    f = [scale_array(mean), scale_array(stdv), scale_array(length)]
    d = np.zeros(len(f[0]), dtype=np.float64)
    d[1:] = f[0][1:] - f[0][:-1]
    f.append(scale_array(d, with_mean=False))
    out = np.empty((len(d), 4 * len(window)), dtype=np.float64)
    for w, pos in enumerate(window):
        for k in range(4):
            out[:, 4 * w + k] = padded_offset(f[k], pos)
    return out


def events_to_features_reads(mean, stdv, length, read_off, window=(-1, 0, 1)):
    """Concatenated reads (the C ABI's layout): read r is events read_off[r]:read_off[r+1]."""
    outs = [events_to_features(mean[a:b], stdv[a:b], length[a:b], window) if b > a else np.zeros((0, 4 * len(window)))
            for a, b in zip(read_off[:-1], read_off[1:])]      # (the reference divides by zero on an empty table)
    return np.concatenate(outs) if outs else np.zeros((0, 4 * len(window)))
