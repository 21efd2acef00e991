#!/usr/bin/env python
"""Throughput of the `.signal` text parse (ntb_parse_signal; the int(x) for x in line.split() of the reference's
read_signal) on one GPU: text bytes/s with the text resident in HBM, HBM roofline on its algorithmic bytes (text read
once + 2 bytes per sample written), and Python's own split/int (what the reference executes) on one core beside it.

This is the body of ``python bench.py --row signal_parse``; the oracle import below is the checker of the timed output.

    python tools/bench_parse.py [--files 256] [--samples 1000000] [--iters 10]
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
    ap.add_argument("--files", type=int, default=256)
    ap.add_argument("--samples", type=int, default=1000000)
    ap.add_argument("--iters", type=int, default=10)
    a = ap.parse_args()
    lib = _lib.load()
    rng = np.random.default_rng(0)
    one = (" ".join(map(str, rng.integers(300, 900, a.samples))) + "\n").encode()     # one read, the shipped format
    import oracle_signal as OS
    t0 = time.perf_counter()
    want = np.asarray([int(x) for x in one.decode().split()])                        # the reference's expression
    cpu_rate = len(one) / (time.perf_counter() - t0)
    assert np.array_equal(want, OS.parse_signal(one))
    total = len(one) * a.files
    text = torch.as_tensor(np.frombuffer(one, dtype=np.uint8).copy()).cuda().repeat(a.files)
    off = np.arange(a.files + 1, dtype=np.int64) * len(one)
    cap = a.samples * a.files
    wb = lib.ntb_parse_signal_workspace_bytes(total, a.files)
    ws = torch.empty(wb, dtype=torch.uint8, device="cuda")
    samples = torch.empty(cap, dtype=torch.int16, device="cuda")
    roff = torch.empty(a.files + 1, dtype=torch.int64, device="cuda")
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    call = lambda: _lib.check(lib.ntb_parse_signal(text.data_ptr(), off.ctypes.data, a.files, total, ws.data_ptr(), wb, cap,  # noqa: E731
                                                   samples.data_ptr(), roff.data_ptr(), status.data_ptr(), st))
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
    same = int(status[0]) == 0 and int(roff[-1]) == cap and np.array_equal(samples[-a.samples:].cpu().numpy(), want)
    peaks_p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm = float(json.load(open(peaks_p)).get("hbm_gbs", 6500.0)) if os.path.exists(peaks_p) else 6500.0
    bytes_alg = total + 2.0 * cap
    gbs = bytes_alg / (ms * 1e-3) / 1e9
    print(json.dumps({"metric": "signal_parse_text_bytes_per_s", "value": total / (ms * 1e-3), "unit": "bytes/s",
                      "files": a.files, "samples_per_file": a.samples, "text_bytes": total, "ms_per_call": ms,
                      "samples_per_s": cap / (ms * 1e-3), "gpu_launches_per_call": 2,
                      "identical_to_python_parse_on_sample": bool(same),
                      "note": "the call synchronises once to upload its chunk table; that wait is inside the timed region",
                      "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s", "frac": gbs / hbm,
                                   "algorithmic_bytes_per_text_byte": bytes_alg / total},
                      "cpu_baseline": {"value": cpu_rate, "unit": "bytes/s", "cores": 1, "kind": "reference",
                                       "sample": "1 file x %d samples, [int(x) for x in text.split()]" % a.samples}}))


if __name__ == "__main__":
    main()
