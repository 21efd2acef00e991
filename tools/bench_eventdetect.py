#!/usr/bin/env python
"""Throughput of event detection (ntb_event_detect; nanonet minknow_event_detect) on one GPU: samples/s with the raw
signal resident in HBM (CUDA events around the five launches), through the host entry point, the HBM roofline fraction
of its algorithmic bytes (8 B of raw signal in per sample + 32 B per event out), and the reference's own compiled C
(oracle/_ref, driven by the NumPy restatement's Python for the peak detector and event table) on one core beside it.

This is the body of ``python bench.py --row event_detect``; the oracle import below is that bench's cpu_baseline leg (the checker
and the timed CPU arm), never part of the measured path.

    python tools/bench_eventdetect.py [--reads 8192] [--samples 50000] [--iters 5]
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]
import nanopore_rnn_b200  # noqa: E402,F401
from nanopore_rnn_b200 import _lib  # noqa: E402


class Detector(C.Structure):           # the reference's struct (filters.h:21-32), to call its compiled C directly
    _fields_ = [("DEF_PEAK_POS", C.c_int), ("DEF_PEAK_VAL", C.c_double), ("signal", C.POINTER(C.c_double)),
                ("signal_length", C.c_size_t), ("threshold", C.c_double), ("window_length", C.c_size_t),
                ("masked_to", C.c_size_t), ("peak_pos", C.c_int), ("peak_value", C.c_double), ("valid_peak", C.c_bool)]


def reference_c_rate(x, OD, reps=5):
    """samples/s of the reference's own C (oracle/_ref: cumulative sums, two t-statistics, peak detector) plus the event
    table in Python as the reference builds it; falls back to the NumPy restatement where oracle/_ref is absent."""
    path = os.path.join(ROOT, "oracle", "_ref", "libnanonetfilters.so")
    if not os.path.exists(path):
        t0 = time.perf_counter()
        OD.minknow_event_detect(x, 4000.0, [3, 6], [1.4, 1.1], 0.2)
        return len(x) / (time.perf_counter() - t0), "port", "NumPy restatement (Python peak detector)"
    ref = C.CDLL(path)
    dp = C.POINTER(C.c_double)
    n = len(x)
    t0 = time.perf_counter()
    for _ in range(reps):
        s, q = np.empty_like(x), np.empty_like(x)
        ref.compute_sum_sumsq(x.ctypes.data_as(dp), s.ctypes.data_as(dp), q.ctypes.data_as(dp), C.c_size_t(n))
        ts = []
        for w in (3, 6):
            t = np.empty_like(x)
            ref.compute_tstat(s.ctypes.data_as(dp), q.ctypes.data_as(dp), t.ctypes.data_as(dp), C.c_size_t(n), C.c_size_t(w),
                              C.c_bool(False))
            ts.append(t)
        dets = [Detector(-1, 1e100, t.ctypes.data_as(dp), n, th, w, 0, -1, 1e100, False)
                for t, th, w in zip(ts, (1.4, 1.1), (3, 6))]
        peaks = np.zeros(n, dtype=np.uintp)
        ref.short_long_peak_detector(C.byref(dets[0]), C.byref(dets[1]), C.c_double(0.2), peaks.ctypes.data_as(C.c_void_p))
        OD.construct_events(s, q, peaks[np.nonzero(peaks)].astype(np.int64), 4000.0)
    return reps * n / (time.perf_counter() - t0), "reference", "filters.c compiled unmodified + the event table in Python"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=8192)
    ap.add_argument("--samples", type=int, default=50000)
    ap.add_argument("--iters", type=int, default=5)
    a = ap.parse_args()
    lib = _lib.load()
    rng = np.random.default_rng(0)
    # one synthetic squiggle, rolled per read (generation cost only; every read differs in phase and noise seed)
    base = np.repeat(rng.uniform(55.0, 125.0, a.samples // 8 + 1), 8)[:a.samples]
    total = a.reads * a.samples
    gen = torch.Generator(device="cuda").manual_seed(1)
    raw = (torch.as_tensor(base).cuda().repeat(a.reads) + 1.2 * torch.randn(total, dtype=torch.float64, device="cuda", generator=gen))
    off = np.arange(a.reads + 1, dtype=np.int64) * a.samples
    wl, th, ph = np.asarray([3, 6], dtype=np.int32), np.asarray([1.4, 1.1]), 0.2
    peaks_p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm = float(json.load(open(peaks_p)).get("hbm_gbs", 6500.0)) if os.path.exists(peaks_p) else 6500.0

    import oracle_eventdetect as OD
    sample = raw[:a.samples].cpu().numpy()
    want, _ = OD.minknow_event_detect(sample, 4000.0, [3, 6], [1.4, 1.1], 0.2)
    cpu_rate, cpu_kind, cpu_what = reference_c_rate(sample, OD)

    wb = lib.ntb_event_detect_workspace_bytes(a.reads, total)
    ws = torch.empty(wb, dtype=torch.uint8, device="cuda")
    cap = total // 2
    ev = torch.empty((cap, 4), dtype=torch.float64, device="cuda")
    eo = torch.empty(a.reads + 1, dtype=torch.int64, device="cuda")
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    call = lambda: _lib.check(lib.ntb_event_detect(raw.data_ptr(), off.ctypes.data, a.reads, total, 4000.0, wl.ctypes.data,  # noqa: E731
                                                   th.ctypes.data, ph, ws.data_ptr(), wb, cap, ev.data_ptr(), eo.data_ptr(),
                                                   status.data_ptr(), st))
    for _ in range(3):
        call()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        call()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.iters
    n_ev = int(eo[-1])
    n0 = int(eo[1])
    got = ev[:n0].cpu().numpy()
    same = int(status[0]) == 0 and n0 == len(want) and np.array_equal(got[:, :3], want[:, :3])
    bytes_alg = 8.0 * total + 32.0 * n_ev
    gbs = bytes_alg / (ms * 1e-3) / 1e9
    print(json.dumps({"metric": "event_detect_samples_per_s", "value": total / (ms * 1e-3), "unit": "samples/s",
                      "reads": a.reads, "samples_per_read": a.samples, "events": n_ev, "ms_per_call": ms,
                      "gpu_launches_per_call": 5, "boundaries_and_means_identical_on_sample": bool(same),
                      "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s", "frac": gbs / hbm,
                                   "algorithmic_bytes_per_sample": bytes_alg / total},
                      "cpu_baseline": {"value": cpu_rate, "unit": "samples/s", "cores": 1, "kind": cpu_kind,
                                       "sample": "1 read x %d samples, %s" % (a.samples, cpu_what)}}))


if __name__ == "__main__":
    main()
