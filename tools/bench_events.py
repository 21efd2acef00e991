#!/usr/bin/env python
"""Throughput of the event-feature kernel (ntb_event_features; nanonet features.py::events_to_features) on one GPU:
events/s device-resident and through the host entry point, the HBM roofline fraction of its algorithmic bytes
(24 B in + 12 x sizeof(out) B out per event), and the NumPy restatement (NumPy's own reductions = what the reference
executes) timed on one host core beside it.  Prints one JSON line per output type.

This is the body of ``python bench.py --row event_features``; the oracle import below is that bench's cpu_baseline leg (the checker
and the timed CPU arm), never part of the measured path.

    python tools/bench_events.py [--reads 2048] [--events 9000] [--iters 10]
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=2048)
    ap.add_argument("--events", type=int, default=9000)      # the reference's max_len (features.py:78)
    ap.add_argument("--iters", type=int, default=10)
    a = ap.parse_args()
    lib = _lib.load()
    rng = np.random.default_rng(0)
    total = a.reads * a.events
    mean = rng.choice(np.linspace(60.0, 120.0, 64), total) + rng.normal(0, 1.5, total)
    stdv = np.abs(rng.normal(1.5, 0.4, total)) + 0.2
    length = rng.integers(3, 60, total) / 4000.0
    off = np.arange(a.reads + 1, dtype=np.int64) * a.events
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    hbm = float(peaks.get("hbm_gbs_sustained", peaks.get("hbm_gbs", 6500.0)) or 6500.0)

    # CPU: the restatement with NumPy's reductions, one core, a bounded sample of the same reads
    import oracle_events as OE
    OE._SUM = np.add.reduce
    n_cpu = min(a.reads, 64)
    t0 = time.perf_counter()
    want = OE.events_to_features_reads(mean[:n_cpu * a.events], stdv[:n_cpu * a.events], length[:n_cpu * a.events],
                                       off[:n_cpu + 1])
    cpu_s = time.perf_counter() - t0
    cpu_rate = n_cpu * a.events / cpu_s

    d = [torch.as_tensor(v).cuda() for v in (mean, stdv, length)]
    st = torch.cuda.current_stream()
    for f64 in (1, 0):
        out = torch.empty((total, 12), dtype=torch.float64 if f64 else torch.float32, device="cuda")
        wb = lib.ntb_event_features_workspace_bytes(a.reads, total)
        ws = torch.empty(wb, dtype=torch.uint8, device="cuda")
        call = lambda: _lib.check(lib.ntb_event_features(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(),  # noqa: E731
                                                         off.ctypes.data, a.reads, total, f64, ws.data_ptr(), wb,
                                                         out.data_ptr(), C.c_void_p(st.cuda_stream)))
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
        got = out[:n_cpu * a.events].cpu().numpy()
        exact = bool(np.array_equal(got, want if f64 else want.astype(np.float32)))
        # host entry point: pageable NumPy buffers in, NumPy buffer out
        ho = np.empty((total, 12), dtype=np.float64 if f64 else np.float32)
        hcall = lambda: _lib.check(lib.ntb_event_features_host(mean.ctypes.data, stdv.ctypes.data,  # noqa: E731
                                                               length.ctypes.data, off.ctypes.data, a.reads, total,
                                                               f64, ho.ctypes.data, None))
        hcall()
        t0 = time.perf_counter()
        hcall()
        host_s = time.perf_counter() - t0
        bytes_ev = 24 + 12 * (8 if f64 else 4)
        gbs = bytes_ev * total / (ms * 1e-3) / 1e9
        print(json.dumps({"metric": "event_features_events_per_s", "value": total / (ms * 1e-3), "unit": "events/s",
                          "out": "f64" if f64 else "f32", "reads": a.reads, "events_per_read": a.events,
                          "ms_per_launch": ms, "bit_identical_to_numpy_on_sample": exact,
                          "e2e_host_buffers_events_per_s": total / host_s,
                          "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s",
                                       "frac": gbs / hbm, "algorithmic_bytes_per_event": bytes_ev},
                          "cpu_baseline": {"value": cpu_rate, "unit": "events/s", "cores": 1, "kind": "port",
                                           "sample": "%d reads x %d events, NumPy reductions" % (n_cpu, a.events)}}))


if __name__ == "__main__":
    main()
