#!/usr/bin/env python
"""Throughput of the consensus assembly (ntb_assemble_reads; process_output / simple_assembly of the reference) on one
GPU: windows/s with the label rows resident in HBM (CUDA events around the five kernels), through the host entry point,
the HBM roofline fraction of its algorithmic bytes (4 T bytes of labels in per window + 1 byte per consensus base out),
and the difflib host mirror (the calls the reference makes) on one core beside it.  Prints one JSON line.

This is the body of ``python bench.py --row assembly``; the oracle import below is that bench's cpu_baseline leg (the checker
and the timed CPU arm), never part of the measured path.

    python tools/bench_assembly.py [--reads 1024] [--windows 1024] [--iters 10]
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
from nanopore_rnn_b200 import _lib, assembly, dataset  # noqa: E402


def job(rng, n_reads, nw, T=100, err=0.05):
    """Windows sliding by ~8 bases over a random sequence per read, ~27 called bases per window, with errors."""
    truth = rng.integers(0, 4, (n_reads, 8 * nw + 40)).astype(np.int32)
    labels = np.full((n_reads, nw, T), -1, dtype=np.int32)
    idx = np.arange(30)
    for w in range(nw):
        seg = truth[:, 8 * w + idx]
        flip = rng.random(seg.shape) < err
        seg = np.where(flip, rng.integers(0, 5, seg.shape), seg)
        keep = int(rng.integers(24, 31))
        labels[:, w, :keep] = seg[:, :keep]
    return labels.reshape(-1, T), np.arange(n_reads + 1, dtype=np.int64) * nw


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=1024)
    ap.add_argument("--windows", type=int, default=1024)
    ap.add_argument("--iters", type=int, default=10)
    a = ap.parse_args()
    lib = _lib.load()
    labels, off = job(np.random.default_rng(0), a.reads, a.windows)
    n, T = labels.shape
    peaks_p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm = float(json.load(open(peaks_p)).get("hbm_gbs", 6500.0)) if os.path.exists(peaks_p) else 6500.0

    n_cpu = min(a.reads, 4)                                   # the difflib mirror: ~1 ms per window pair
    t0 = time.perf_counter()
    want = [dataset.assemble_read([labels[x:y]]) for x, y in zip(off[:n_cpu], off[1:n_cpu + 1])]
    cpu_rate = int(off[n_cpu]) / (time.perf_counter() - t0)

    cap = int((labels != -1).sum())
    d = torch.as_tensor(labels).cuda()
    ws_b = lib.ntb_assemble_workspace_bytes(n, a.reads, T, cap)
    ws = torch.empty(ws_b, dtype=torch.uint8, device="cuda")
    cons = torch.empty(cap, dtype=torch.uint8, device="cuda")
    coff = torch.empty(a.reads + 1, dtype=torch.int64, device="cuda")
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    call = lambda: _lib.check(lib.ntb_assemble_reads(d.data_ptr(), off.ctypes.data, a.reads, n, T, cap, ws.data_ptr(),  # noqa: E731
                                                     ws_b, cons.data_ptr(), coff.data_ptr(), None, status.data_ptr(), st))
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
    co = coff.cpu().numpy()
    text = cons[:co[-1]].cpu().numpy().tobytes().decode()
    same = [text[x:y] for x, y in zip(co[:n_cpu], co[1:n_cpu + 1])] == want and int(status[0]) == 0
    assembly.assemble_reads(labels, off)
    t0 = time.perf_counter()
    assembly.assemble_reads(labels, off)
    host_s = time.perf_counter() - t0
    bytes_alg = 4.0 * T * n + float(co[-1])
    gbs = bytes_alg / (ms * 1e-3) / 1e9
    print(json.dumps({"metric": "assembly_windows_per_s", "value": n / (ms * 1e-3), "unit": "windows/s",
                      "reads": a.reads, "windows_per_read": a.windows, "called_bases": cap, "consensus_bases": int(co[-1]),
                      "ms_per_call": ms, "gpu_launches_per_call": 7, "identical_to_difflib_mirror_on_sample": bool(same),
                      "e2e_host_buffers_windows_per_s": n / host_s,
                      "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s", "frac": gbs / hbm,
                                   "algorithmic_bytes_per_window": bytes_alg / n},
                      "cpu_baseline": {"value": cpu_rate, "unit": "windows/s", "cores": 1, "kind": "port",
                                       "sample": "%d reads x %d windows, difflib host mirror" % (n_cpu, a.windows)}}))


if __name__ == "__main__":
    main()
