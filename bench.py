#!/usr/bin/env python
"""bench.py -- signal samples/sec basecalled (NanoTensor BLSTM forward) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--windows M]

A *step* is one pass of the hot path (conv/BN front-end, BLSTM projections +
recurrences, fused chiron_fnn/dense/CTC-greedy readout) over one batch of synthetic windows of
BASELINE.json configs[1]: the ``run_nanotensor.config.json`` network (3 residual(256) +
3 BLSTM(100, forget_bias 5) + chiron_fnn, CtcLoss C=4, T=100), 1 048 576 windows in BN groups
of 10 (the config's batch_size), the last window of every 1000-window read ragged.  Weak scaling:
every rank processes its own ``--windows`` windows (reads shard across GPUs, no collective on
the data path); ``value`` = valid samples of all ranks / max-over-ranks device time.

One JSON line is printed by rank 0 (keys: see DESIGN.md "Measurement").  ``--impl reference``
times the CPU restatement of the reference's TensorFlow forward (oracle/oracle_cpu.cpp, C++/OpenMP, all host
threads) on a bounded sample of the same workload: TensorFlow 1.4 itself is unavailable on the
box.  ``--row event_features|assembly|event_detect|signal_parse`` measures one of the rows the path was
widened into with that row's own metric instead (tools/bench_*.py).
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ROWS = {"event_features": "bench_events.py", "assembly": "bench_assembly.py", "event_detect": "bench_eventdetect.py",
        "signal_parse": "bench_parse.py"}
METRIC = "signal samples/sec basecalled (BLSTM fwd)"
UNIT = "samples/s"
WORKLOAD = "run_nanotensor.config.json network, synthetic windows"
T, BATCH, NCLASS, WIN_PER_READ = 100, 10, 4, 1000
# algorithmic FLOPs per window (multiply-add = 2)
FLOPS_CONV, FLOPS_PROJ, FLOPS_REC, FLOPS_READOUT = 209.8e6, 105.0e6, 48.0e6, 0.12e6


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--windows", type=int, default=1048576, help="windows per GPU per step")
    ap.add_argument("--cpu-windows", type=int, default=0,
                    help="windows in the CPU baseline sample (0 = auto-size to ~10-20 s)")
    ap.add_argument("--chunk-windows", type=int, default=0,
                    help="windows per forward call (0 = as many as the workspace budget allows)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true", help="skip the end-to-end (host buffers) leg")
    ap.add_argument("--row", default="forward", choices=["forward"] + sorted(ROWS),
                    help="which SURVEY section-8 row to measure: the BLSTM forward (default, the contract's line) or one "
                         "of the widened rows (f): their own metric, roofline and cpu_baseline, one GPU")
    args, rest = ap.parse_known_args()
    args.rest = rest
    return args


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "src": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "src": "fallback"}


# ---------------------------------------------------------------------------- workload
def make_workload(n_windows, seed):
    """Windows of synthetic z-scored reads + BN group offsets."""
    from nanopore_rnn_b200 import engine, synthetic as syn
    n_reads = max(1, -(-n_windows // WIN_PER_READ))
    wpr = WIN_PER_READ if n_windows >= WIN_PER_READ else n_windows
    x, lens, _ = syn.synthetic_reads(n_reads, wpr, T, seed=seed)
    x, lens = x[:n_windows], lens[:n_windows]
    counts = [wpr] * n_reads
    counts[-1] -= n_reads * wpr - n_windows
    goff = engine.group_offsets(counts, BATCH)
    return x, lens, goff


def network_and_weights():
    from nanopore_rnn_b200 import synthetic as syn
    net = syn.NETWORKS["run_nanotensor"]
    return net, syn.random_weights(net, 1, NCLASS, seed=2)


# ---------------------------------------------------------------------------- clocks
class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons of one GPU through NVML while the timed region runs."""

    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown",
               0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown"}

    def __init__(self, index, period=0.1):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.sm, self.reasons, self.sm_max = [], set(), None
        self._stop_evt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.sm_max = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception as e:          # NVML missing: report it instead of inventing clocks
            self.nv, self.err = None, repr(e)

    def run(self):
        if self.nv is None:
            return
        while not self._stop_evt.is_set():
            try:
                self.sm.append(int(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
                mask = int(self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop_evt.wait(self.period)

    def finish(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join(timeout=2)
        if self.nv is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "error": getattr(self, "err", "nvml")}
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.sm_max,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


# ---------------------------------------------------------------------------- CPU baseline
def tf_probe():
    """Run-time probe for the real reference arithmetic (oracle/tf_probe.py): `import tensorflow` is TRIED on the box
    this runs on and the outcome recorded in the JSON line; if it ever imports, the goldens are re-derived with TF ops
    and the distance to the committed goldens is reported too (that would pin the CPU restatements)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import tf_probe as tp
    out = tp.probe()
    out["observed"] = tp.describe()
    if out["importable"]:
        try:
            out["golden_max_abs_err"] = tp.pin_against_goldens(os.path.join(ROOT, "tests", "golden"))
        except Exception as e:
            out["golden_error"] = repr(e)[:200]
    return out


def _cpu_forward():
    """The CPU arm: oracle/oracle_cpu.cpp (C++/OpenMP restatement of the reference's TF forward, all host threads)
    -- the faster of the two CPU restatements; oracle/oracle_torch.py if the library has not been built."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    net, w = network_and_weights()
    cores = os.cpu_count() or 1
    try:
        import oracle_cpu as Cc
        Cc.load()

        def run(x, lens, goff):
            Cc.forward(x, lens, net, w, "CtcLoss", group_off=goff, n_threads=cores)   # logits + CTC-greedy labels
        return run, cores, "C++/OpenMP restatement (oracle/oracle_cpu.cpp)"
    except Exception:
        import torch
        import oracle_np as A
        import oracle_torch as B
        torch.set_num_threads(cores)

        def run(x, lens, goff):
            A.ctc_greedy(B.forward_job(x, lens, net, w, "CtcLoss", BATCH), lens)
        return run, cores, "torch-CPU restatement (oracle/oracle_torch.py)"


def _cpu_sample_size(run, target_s, cap):
    x, lens, goff = make_workload(BATCH * 64, 1234)
    run(x[:BATCH * 8], lens[:BATCH * 8], goff[:9])           # warm the thread pool / allocators
    t0 = time.perf_counter()
    run(x, lens, goff)
    dt = time.perf_counter() - t0
    return int(max(BATCH * 64, min(cap, (target_s / dt) * BATCH * 64)) // BATCH * BATCH)


def cpu_baseline(n_windows):
    """Times the CPU restatement of the reference's forward (per-batch BN kept, BN groups of 10) on a bounded sample."""
    run, cores, what = _cpu_forward()
    if n_windows <= 0:
        n_windows = _cpu_sample_size(run, 12.0, 65536)
    x, lens, goff = make_workload(n_windows, 1234)
    t0 = time.perf_counter()
    run(x, lens, goff)
    dt = time.perf_counter() - t0
    tfp = tf_probe()
    return {"value": int(lens.sum()) / dt, "unit": UNIT, "cores": cores, "kind": "port", "tf_probe": tfp,
            "sample": "%d windows (T=%d, BN groups of %d) of the same workload, %.1f s; %s, %d threads; %s"
                      % (n_windows, T, BATCH, dt, what, cores, tfp["observed"])}


# ---------------------------------------------------------------------------- reference arm
def run_reference(args):
    """--impl reference: the CPU restatement on the host cores; each step is a bounded sample of the
    workload (a few seconds), rank 0 only."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return 0
    os.environ.pop("OMP_NUM_THREADS", None)       # torchrun pins it to 1; the arm uses every host thread explicitly
    run, cores, what = _cpu_forward()
    nw = args.cpu_windows
    if nw <= 0:
        nw = _cpu_sample_size(run, 5.0, 32768)
    x, lens, goff = make_workload(nw, 1234)
    samples = int(lens.sum())
    for _ in range(args.warmup):
        run(x, lens, goff)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        run(x, lens, goff)
    dt = (time.perf_counter() - t0) / max(args.steps, 1)
    value = samples / dt
    tfp = tf_probe()
    sample = ("%d windows per step (T=%d, BN groups of %d), %s, %d threads; %s"
              % (nw, T, BATCH, what, cores, tfp["observed"]))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "windows_per_step": nw, "seq_len": T, "bn_group": BATCH,
                       "n_classes": NCLASS},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                             "tf_probe": tfp},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------- B200 arm
def run_b200(args):
    import torch
    import nanopore_rnn_b200  # noqa: F401
    from nanopore_rnn_b200 import engine
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(v):
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # each rank owns its own reads (weak scaling): no collective on the data path
    x, lens, goff = make_workload(args.windows, seed=100 + rank)
    n = len(x)
    samples = int(lens.sum())
    net, w = network_and_weights()
    m = engine.Model(net, w, "CtcLoss", T, 1, NCLASS, device=str(dev))

    # ---- device-resident job: inputs already in HBM when the timed region starts
    x_pin = torch.from_numpy(x).pin_memory()
    lens_pin = torch.from_numpy(lens).pin_memory()
    x_d, lens_d = x_pin.to(dev), lens_pin.to(dev)
    chunks = engine.plan_chunks(goff, args.chunk_windows or m.max_chunk_windows(), m.chunk_align())
    chunk_goff = [torch.from_numpy(np.ascontiguousarray(goff[g:h + 1] - goff[g], dtype=np.int32)).to(dev)
                  for g, h in chunks]
    labels_d = torch.empty(n * T, dtype=torch.int32, device=dev)
    llen_d = torch.empty(n, dtype=torch.int32, device=dev)
    launches = [0]

    def step():
        for (g, h), go in zip(chunks, chunk_goff):
            w0, w1 = int(goff[g]), int(goff[h])
            m.forward_device(x_d[w0:w1], lens_d[w0:w1], go, labels_out=labels_d[w0 * T:w1 * T],
                             label_len_out=llen_d[w0:w1])
            launches[0] += m.last_launches()

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    launches[0] = 0
    m.set_profile(True)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    barrier()
    clocks = sampler.finish()
    ms_total = max_over_ranks(ev0.elapsed_time(ev1))
    stage = m.stage_ms()
    m.set_profile(False)
    n_launch = launches[0]
    ms_step = ms_total / max(args.steps, 1)
    total_samples = sum_over_ranks(samples)
    value = total_samples / (ms_step * 1e-3)

    # ---- end to end through the public API with HOST buffers (pinned), copies inside
    e2e = None
    if not args.no_e2e:
        xh, lh = x_pin.numpy(), lens_pin.numpy()
        for _ in range(2):                                # warm: workspace, staging, and BOTH pinned result blocks
            labels, llen, _ = m.forward_host(xh, lh, goff)    #  (the previous result is still alive while a call runs)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            labels, llen, _ = m.forward_host(xh, lh, goff)
        torch.cuda.synchronize()
        dt = max_over_ranks(time.perf_counter() - t0) / max(args.steps, 1)
        e2e = {"value": total_samples / dt, "unit": UNIT,
               "h2d_bytes_per_step": int(sum_over_ranks(x.nbytes + lens.nbytes + goff.nbytes)),
               "d2h_bytes_per_step": int(sum_over_ranks(labels.nbytes + llen.nbytes)), "ms_per_step": dt * 1e3}

    # ---- roofline of the dominant kernel: k_gemm_tc<2>, the tcgen05 GEMM of the LSTM input projections (three
    # launches per forward, the largest share of the launch list in profiles/).  Its launches are bracketed by CUDA
    # events on the forward's stream (stage 4 of ntb_model_stage_ms); achieved = SURVEY.md 8(d)'s algorithmic projection
    # FLOPs per window x windows / that time.  The per-stage table (same events) covers the rest of the step.
    pk = peaks()
    proj_ms = stage.get("projection_gemm", 0.0)
    stage4 = {"conv_bn": stage.get("conv_bn", 0.0),
              "projection": stage.get("projection_split", 0.0) + proj_ms,
              "recurrence": stage.get("recurrence", 0.0), "readout": stage.get("readout", 0.0)}
    flops = {"conv_bn": FLOPS_CONV, "projection": FLOPS_PROJ, "recurrence": FLOPS_REC, "readout": FLOPS_READOUT}
    roof = None
    if proj_ms > 0:
        n_launch_k = 3 * len(chunks) * args.steps
        achieved = FLOPS_PROJ * n * args.steps / (proj_ms * 1e-3) / 1e12
        peak = pk["bf16_tflops_sustained"]
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            per_window = json.load(open(tp)).get("k_gemm_tc_xproj_dram_bytes_per_window")
            if per_window:
                traffic = per_window * n / len(chunks)      # bytes per launch at this chunk size (ncu: --set full)
        roof = {"bound": "tensor", "kernel": "k_gemm_tc<2> (LSTM input projection GEMM, tcgen05 fp16x3)",
                "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak, "traffic": traffic,
                "peak_source": pk["src"] + " (sustained bf16)",
                "us_per_launch": proj_ms * 1e3 / n_launch_k, "launches": n_launch_k,
                "algorithmic_flops_per_launch": FLOPS_PROJ * n / len(chunks) / 3,
                "tensor_pipe_frac_with_3_passes": 3 * achieved / peak,
                "stage_ms_per_step": {k: v / max(args.steps, 1) for k, v in stage4.items()},
                "stage_tflops": {k: (flops[k] * n * args.steps / (v * 1e-3) / 1e12 if v > 0 else None)
                                 for k, v in stage4.items()},
                "hbm_peak_gbs": pk["hbm_gbs"]}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_baseline(args.cpu_windows)
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": WORKLOAD, "windows_per_gpu": n, "seq_len": T, "bn_group": BATCH,
                           "n_classes": NCLASS, "chunks_per_step": len(chunks),
                           "l2": "inputs (%.0f MB) and activations larger than L2" % (x.nbytes / 1e6),
                           "parallelism": "reads sharded over %d GPU(s), no collective" % world},
                "clocks": clocks, "e2e": e2e, "gpu_launches": int(n_launch), "roofline": roof,
                "cpu_baseline": cpu}
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def run_row(args):
    """The rows SURVEY section 8(f) widened into (event detection, event features, consensus assembly): each has its own
    bench under tools/ -- metric, HBM roofline on its algorithmic bytes, and a cpu_baseline leg (the only place they
    touch oracle/) -- run here so that bench.py stays the one measurement entry point.  Extra flags go through
    (e.g. ``python bench.py --row event_features --reads 2048 --events 9000``)."""
    import runpy
    sys.argv = [os.path.join(ROOT, "tools", ROWS[args.row])] + args.rest
    runpy.run_path(sys.argv[0], run_name="__main__")
    return 0


def main():
    args = parse()
    if args.row != "forward":
        return run_row(args)
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
