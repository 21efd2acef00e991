#!/usr/bin/env python
"""Generates tests/golden/event_detect.npz by RUNNING THE REFERENCE'S OWN CODE (build container only).

SURVEY.md section 8(f) rank 4, second half: ``minknow_event_detect`` (nanonet/nanonet/eventdetection/filters.py:299-322)
= compute_sum_sumsq + compute_tstat (two windows) + short_long_peak_detector (all three in the reference's C file
filters.c, compiled unmodified by oracle/build_ref.sh into oracle/_ref/libnanonetfilters.so) + ``_construct_events``
(filters.py:146-204).  The Python wrappers are cut out of filters.py with ``ast`` and executed against that library
(the module itself imports nanonet.util to locate its .so).  The vectors are therefore the reference's outputs: PINNED.

Inputs: synthetic squiggles in pA (piecewise-constant levels + Gaussian noise, non-integer float64), run with nanonet's
basecalling parameters (features.py:78: windows [3, 6], thresholds [1.4, 1.1], peak height 0.2) and with the function's
defaults ([16, 40], [8.0, 4.0], 1.0); plus reads shorter than twice the window (no t-statistic, one event) and a flat
read.
"""
import ast
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference/nanonet/nanonet/eventdetection/filters.py"
PARAMS = [([3, 6], [1.4, 1.1], 0.2), ([16, 40], [8.0, 4.0], 1.0)]


def reference_namespace():
    from ctypes import POINTER, Structure, c_bool, c_double, c_int, c_size_t
    from numpy.ctypeslib import ndpointer
    lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libnanonetfilters.so"))
    tree = ast.parse(open(REF).read())
    picked = [n for n in tree.body if isinstance(n, (ast.FunctionDef, ast.ClassDef))]
    ns = dict(np=np, ndpointer=ndpointer, c_double=c_double, c_bool=c_bool, c_size_t=c_size_t, c_int=c_int,
              Structure=Structure, POINTER=POINTER, nanonetfilters=lib)
    exec(compile(ast.Module(body=picked, type_ignores=[]), REF, "exec"), ns)
    return ns


def squiggle(rng, n, mean_dwell=9.0, noise=1.2):
    out = np.empty(n)
    i = 0
    while i < n:
        d = 1 + int(rng.geometric(1.0 / mean_dwell))
        out[i:i + d] = rng.uniform(55.0, 125.0)
        i += d
    return out + rng.normal(0.0, noise, n)


def main():
    ns = reference_namespace()
    rng = np.random.default_rng(7)
    raws = [squiggle(rng, 50000), squiggle(rng, 20000, mean_dwell=25.0, noise=2.5), squiggle(rng, 11),
            squiggle(rng, 5), np.full(300, 80.0) + 0.0, squiggle(rng, 1), squiggle(rng, 79), squiggle(rng, 80)]
    out = dict(raw=np.concatenate(raws), read_off=np.cumsum([0] + [len(r) for r in raws]).astype(np.int64),
               sample_rate=np.float64(4000.0))
    for p, (wl, th, ph) in enumerate(PARAMS):
        ev_all, ev_off, pk_all, pk_off = [], [0], [], [0]
        for raw in raws:
            ev = ns["minknow_event_detect"](raw.copy(), 4000.0, window_lengths=wl, thresholds=th, peak_height=ph)
            pk = ns["minknow_event_detect"](raw.copy(), 4000.0, window_lengths=wl, thresholds=th, peak_height=ph,
                                            get_peaks=True)
            ev_all.append(np.stack([ev["start"], ev["length"], ev["mean"], ev["stdv"]], axis=1))
            ev_off.append(ev_off[-1] + len(ev))
            pk_all.append(np.asarray(pk, dtype=np.int64))
            pk_off.append(pk_off[-1] + len(pk))
        out["events_%d" % p] = np.concatenate(ev_all)
        out["event_off_%d" % p] = np.asarray(ev_off, dtype=np.int64)
        out["peaks_%d" % p] = np.concatenate(pk_all)
        out["peak_off_%d" % p] = np.asarray(pk_off, dtype=np.int64)
        out["params_%d" % p] = np.asarray(wl + th + [ph], dtype=np.float64)
        print("params", wl, th, ph, "events per read", np.diff(ev_off))
    np.savez_compressed(os.path.join(HERE, "event_detect.npz"), **out)


if __name__ == "__main__":
    main()
