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
WIN_PER_READ = 1000
# --network: the three networks BASELINE.json's configs name.  The contract line (no flag) is configs[1]; the others
# are extra throughput + roofline lines (profiles/), each parity-tested in tests/test_gpu_parity.py.
WORKLOADS = {
    "run_nanotensor": {"graph": "CtcLoss", "T": 100, "batch": 10, "n_classes": 4, "len_x": 1, "windows": 1048576,
                       "weights_seed": 2, "what": "BASELINE configs[1]: run_nanotensor.config.json network "
                                                  "(3 residual(256) + 3 BLSTM(100) + chiron_fnn, CTC C=4), synthetic windows"},
    "signal_CRNN": {"graph": "CtcLoss", "T": 100, "batch": 20, "n_classes": 6, "len_x": 1, "windows": 1048576,
                    "weights_seed": 12, "what": "BASELINE configs[2]: signal_CRNN.config.json network as shipped "
                                                "(2 residual(256) + 1 BLSTM(100) + chiron_fnn, CTC C=6), synthetic windows"},
    "stress_blstm256": {"graph": "CrossEntropy", "T": 100, "batch": 0, "n_classes": 1025, "len_x": 12, "windows": 262144,
                        "weights_seed": 32, "what": "BASELINE configs[4]: 5 x BLSTM(256), CrossEntropy C=1025, len_x 12, "
                                                    "synthetic N(0,1) windows"},
    # SURVEY.md 8(d), the Chiron shape: chiron/chiron/entry.py:32-33 (segment_len 300, batch_size 1100); same layer stack,
    # 5 classes (4 bases + blank); BN moments over a batch of 1100 segments (330 000 rows per group)
    "chiron_T300": {"graph": "CtcLoss", "T": 300, "batch": 1100, "n_classes": 5, "len_x": 1, "windows": 349800,
                    "weights_seed": 301, "net": "run_nanotensor",
                    "what": "Chiron shape (SURVEY 8d): 3 residual(256) + 3 BLSTM(100) + chiron_fnn, T = 300, CTC C=5, BN "
                            "groups of 1100 windows, synthetic windows (use --step 300 for non-overlapping windows)"},
}
WL = dict(WORKLOADS["run_nanotensor"], name="run_nanotensor")
T, BATCH, NCLASS = WL["T"], WL["batch"], WL["n_classes"]
WORKLOAD = WL["what"]


def select_workload(name):
    global WL, T, BATCH, NCLASS, WORKLOAD
    WL = dict(WORKLOADS[name], name=name)
    T, BATCH, NCLASS, WORKLOAD = WL["T"], WL["batch"], WL["n_classes"], WL["what"]


def flops_per_window(net, T, len_x, n_classes):
    """Algorithmic FLOPs per window by stage (multiply-add = 2): residual layer
    2T(2 Cin Cout + 3 Cout^2 + Cout^2); blstm projection 16 T D H, recurrence 16 T H^2; fnn 2T 2H; dense 2T F C."""
    f = {"conv_bn": 0.0, "projection": 0.0, "recurrence": 0.0, "readout": 0.0}
    d = len_x
    for l in net:
        if l["type"] == "residual_layer":
            c = l["out_channel"]
            f["conv_bn"] += 2.0 * T * (2 * d * c + 4 * c * c)
            d = c
        elif l["type"] == "blstm":
            h = l["size"]
            f["projection"] += 16.0 * T * d * h
            f["recurrence"] += 16.0 * T * h * h
            d = 2 * h
        elif l["type"] == "chiron_fnn":
            f["readout"] += 2.0 * T * d
            d //= 2
    f["readout"] += 2.0 * T * d * n_classes
    return f


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--network", default="run_nanotensor", choices=sorted(WORKLOADS),
                    help="which BASELINE.json network to time (default: configs[1], the contract's line)")
    ap.add_argument("--windows", type=int, default=0, help="windows per GPU per step (0 = the workload's own size)")
    ap.add_argument("--cpu-windows", type=int, default=0,
                    help="windows in the CPU baseline sample (0 = auto-size to ~10-20 s)")
    ap.add_argument("--chunk-windows", type=int, default=0,
                    help="windows per forward call (0 = as many as the workspace budget allows)")
    ap.add_argument("--workload", default="reads", choices=["reads", "windows"],
                    help="reads (default): ONE variable-length read set of --windows windows in total, LPT-sharded over "
                         "the ranks (strong scaling; raw int16 samples -> called reads, gather to rank 0 in e2e); "
                         "windows: every rank its own pre-windowed uniform windows (weak scaling, round 1's line)")
    ap.add_argument("--step", type=int, default=100,
                    help="window step of the reads workload (100 = dense; 300 = the reference's default, dataset.py:337)")
    ap.add_argument("--no-replica", action="store_true", help="reads workload: skip the replica_weak second key")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true", help="skip the end-to-end (host buffers) leg")
    ap.add_argument("--row", default="forward", choices=["forward"] + sorted(ROWS),
                    help="which SURVEY section-8 row to measure: the BLSTM forward (default, the contract's line) or one "
                         "of the widened rows (f): their own metric, roofline and cpu_baseline, one GPU")
    args, rest = ap.parse_known_args()
    args.rest = rest
    select_workload(args.network)
    if args.windows <= 0:
        args.windows = WL["windows"]
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
    """Windows of synthetic z-scored reads + BN group offsets (networks with residual layers;
    the pure-BLSTM stress network has no cross-window coupling: ragged N(0,1) windows, group table = None)."""
    from nanopore_rnn_b200 import engine, synthetic as syn
    if WL["batch"] == 0:
        rng = np.random.default_rng(seed)
        x = rng.standard_normal((n_windows, T, WL["len_x"]), dtype=np.float32)
        lens = np.full(n_windows, T, dtype=np.int32)
        lens[WIN_PER_READ - 1::WIN_PER_READ] = rng.integers(1, T + 1, size=len(lens[WIN_PER_READ - 1::WIN_PER_READ]))
        x *= (np.arange(T)[None, :] < lens[:, None])[:, :, None]
        return x, lens, None
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
    net = syn.NETWORKS[WL.get("net", WL["name"])]
    return net, syn.random_weights(net, WL["len_x"], NCLASS, seed=WL["weights_seed"], graph=WL["graph"])


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

    def sample_once(self):
        """One sample from the calling thread (the end-to-end leg: a polling thread's NVML queries were measured to delay the
        host's wake-up from the stream synchronisation by 0.1-0.9 s now and then, tools/hic.sh)."""
        if self.nv is None:
            return
        try:
            self.sm.append(int(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
            mask = int(self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
            for bit, name in self.REASONS.items():
                if mask & bit:
                    self.reasons.add(name)
        except Exception:
            pass

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
            Cc.forward(x, lens, net, w, WL["graph"], group_off=goff, n_threads=cores)   # logits + CTC-greedy labels
        return run, cores, "C++/OpenMP restatement (oracle/oracle_cpu.cpp)"
    except Exception:
        import torch
        import oracle_np as A
        import oracle_torch as B
        torch.set_num_threads(cores)

        def run(x, lens, goff):
            lg = B.forward_job(x, lens, net, w, WL["graph"], BATCH or len(x))
            A.ctc_greedy(lg, lens) if WL["graph"] == "CtcLoss" else A.argmax_first(lg.reshape(-1, NCLASS))
        return run, cores, "torch-CPU restatement (oracle/oracle_torch.py)"


def _cpu_sample_size(run, target_s, cap):
    b = BATCH or 16
    x, lens, goff = make_workload(b * 64, 1234)
    run(x[:b * 8], lens[:b * 8], goff[:9] if goff is not None else None)     # warm the thread pool / allocators
    t0 = time.perf_counter()
    run(x, lens, goff)
    dt = time.perf_counter() - t0
    return int(max(b * 64, min(cap, (target_s / dt) * b * 64)) // b * b)


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
            # same key as the B200 arm's default line (one read set over N GPUs); the CPU arm itself always runs on rank 0
            "scaling": "strong" if (args.workload == "reads" and WL["graph"] == "CtcLoss" and WL["len_x"] == 1) else "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD + "; CPU arm: a bounded sample of windows of the same network per step, forward "
                                   "only (the B200 arm's job adds z-score / windowing and the consensus assembly)",
                       "windows_per_step": nw, "seq_len": T, "bn_group": BATCH, "n_classes": NCLASS},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                             "tf_probe": tfp},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------- B200 arm
class Ctx(object):
    """Per-process plumbing: device, NCCL group (torch.distributed), the model, rank-reduction helpers."""

    def __init__(self):
        import torch
        import nanopore_rnn_b200  # noqa: F401
        from nanopore_rnn_b200 import engine
        self.torch = torch
        self.rank = int(os.environ.get("RANK", 0))
        self.world = int(os.environ.get("WORLD_SIZE", 1))
        self.local = int(os.environ.get("LOCAL_RANK", 0))
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.net, self.w = network_and_weights()
        self.m = engine.Model(self.net, self.w, WL["graph"], T, WL["len_x"], NCLASS, device=str(self.dev))

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def _reduce(self, v, op):
        if self.dist is None:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=op)
        return float(t.item())

    def max_over_ranks(self, v):
        return self._reduce(v, self.dist.ReduceOp.MAX if self.dist else None)

    def sum_over_ranks(self, v):
        return self._reduce(v, self.dist.ReduceOp.SUM if self.dist else None)

    def all_values(self, v):
        if self.dist is None:
            return [v]
        t = self.torch.zeros(self.world, dtype=self.torch.float64, device=self.dev)
        t[self.rank] = v
        self.dist.all_reduce(t)
        return [float(a) for a in t.tolist()]

    def close(self):
        if self.dist is not None:
            self.dist.barrier()
            self.dist.destroy_process_group()


def roofline_of(cx, stage, n, n_chunks, steps, ms_step):
    """roofline / roofline_stages from the stage marks of ``steps`` forwards over ``n`` windows in ``n_chunks`` chunks.
    Dominant kernel: k_gemm_tc<2>, the persistent tcgen05 GEMM -- in the CRNN networks its largest role is the LSTM
    input projection (one launch per BLSTM layer and chunk), bracketed by its own CUDA events on the forward's stream
    (stage 4 of ntb_model_stage_ms); in the pure-BLSTM stress network it runs once per timestep and direction pair
    inside the recurrence stage, so the stage marks are the kernel's.  achieved = SURVEY.md 8(d)'s algorithmic FLOPs
    per window x windows / that time."""
    net = cx.net
    pk = peaks()
    proj_ms = stage.get("projection_gemm", 0.0)
    stage4 = {"conv_bn": stage.get("conv_bn", 0.0),
              "projection": stage.get("projection_split", 0.0) + proj_ms,
              "recurrence": stage.get("recurrence", 0.0), "readout": stage.get("readout", 0.0)}
    flops = flops_per_window(net, T, WL["len_x"], NCLASS)
    peak = pk["bf16_tflops_sustained"]
    n_blstm = sum(1 for l in net if l["type"] == "blstm")
    stages = {}
    for k, v in stage4.items():
        tf_s = flops[k] * n * steps / (v * 1e-3) / 1e12 if v > 0 else None
        stages[k] = {"ms_per_step": v / steps, "algorithmic_mflop_per_window": flops[k] / 1e6, "achieved_tflops": tf_s,
                     "frac": (tf_s / peak if tf_s else None),
                     "tensor_pipe_frac_with_3_passes": (3 * tf_s / peak if tf_s else None)}
    stages["readout"]["bound"] = "hbm"          # reads the last layer's activations once; FLOPs are negligible
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath) and WL["name"] == "run_nanotensor" and stage4["conv_bn"] > 0:
        # the conv/BN stage is co-limited: its kernels move 2.9 MB of DRAM traffic per window (ncu, profiles/) next to the
        # three-pass tensor work -- both fractions are reported, they add up rather than overlap
        bpw = json.load(open(tpath)).get("conv_bn_stage_dram_bytes_per_window")
        if bpw:
            gbs = bpw * n * steps / (stage4["conv_bn"] * 1e-3) / 1e9
            stages["conv_bn"].update({"dram_mb_per_window": bpw / 1e6, "achieved_gbs": gbs, "hbm_frac": gbs / pk["hbm_gbs"]})
    roof = None
    if proj_ms > 0:
        kname, k_ms, k_flops = "k_gemm_tc<2> (LSTM input projection GEMM, tcgen05 fp16x3)", proj_ms, flops["projection"]
        n_launch_k = n_blstm * n_chunks * steps
        tkey = "k_gemm_tc_xproj_dram_bytes_per_window"
    elif stage4["recurrence"] > 0 and all(l["size"] <= 100 for l in net if l["type"] == "blstm"):
        # the fused BLSTM kernel: input projection AND recurrence of a layer in one persistent CTA-pair launch, so the
        # "recurrence" stage marks bracket exactly its launches and its FLOPs are projection + recurrence
        kname = "k_blstm_fused (x_t.W_ih + h_{t-1}.W_hh + cell update per timestep, tcgen05 cta_group::2 fp16x3)"
        k_ms, k_flops = stage4["recurrence"], flops["projection"] + flops["recurrence"]
        n_launch_k = n_blstm * n_chunks * steps
        tkey = "k_blstm_fused_dram_bytes_per_window"
        fused_tf = k_flops * n * steps / (k_ms * 1e-3) / 1e12
        stages["projection"] = {"ms_per_step": stage4["projection"] / steps, "fused_into": "recurrence",
                                "what": "only the operand split of a BLSTM layer whose input is not already in operand planes"}
        stages["recurrence"].update({"includes": "projection (fused kernel)", "algorithmic_mflop_per_window": k_flops / 1e6,
                                     "achieved_tflops": fused_tf, "frac": fused_tf / peak,
                                     "tensor_pipe_frac_with_3_passes": 3 * fused_tf / peak})
    elif stage4["recurrence"] > 0:
        # H > 100: one tcgen05 GEMM launch per timestep; the projection x_t.W_ih is part of that GEMM's contraction
        # ([W_ih ; W_hh] along K, H % 64 == 0), so its FLOPs belong to this kernel
        kname = "k_gemm_tc<2> (per-timestep GEMM [x_t | h_{t-1}].[W_ih ; W_hh] with fused cell epilogue, tcgen05 fp16x3)"
        inside, d = 0.0, WL["len_x"]
        for l in net:
            if l["type"] == "blstm":
                if d % 4 == 0 and l["size"] % 64 == 0:          # narrow inputs are zero-padded to one 64-wide k-block
                    inside += 16.0 * T * d * l["size"]
                d = 2 * l["size"]
        k_ms, k_flops = stage4["recurrence"], flops["recurrence"] + inside
        n_launch_k = n_blstm * T * n_chunks * steps
        tkey = "k_gemm_tc_step_dram_bytes_per_window"
        step_tf = k_flops * n * steps / (k_ms * 1e-3) / 1e12
        stages["recurrence"].update({"includes": "the input projections (same GEMM)",
                                     "algorithmic_mflop_per_window": k_flops / 1e6, "achieved_tflops": step_tf,
                                     "frac": step_tf / peak, "tensor_pipe_frac_with_3_passes": 3 * step_tf / peak})
        rest = flops["projection"] - inside
        if stage4["projection"] > 0:
            ptf = rest * n * steps / (stage4["projection"] * 1e-3) / 1e12
            stages["projection"].update({"what": "operand splits of the layer inputs (and any projection left outside the step GEMM)",
                                         "algorithmic_mflop_per_window": rest / 1e6, "achieved_tflops": ptf, "frac": ptf / peak,
                                         "tensor_pipe_frac_with_3_passes": 3 * ptf / peak})
    else:
        kname = None
    if kname:
        achieved = k_flops * n * steps / (k_ms * 1e-3) / 1e12
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp) and WL["name"] != "signal_CRNN":
            per_window = json.load(open(tp)).get(tkey)
            if per_window:     # bytes per launch at this chunk size (ncu --set full, per window, profiles/)
                traffic = per_window * n / n_chunks * (1.0 if (proj_ms > 0 or "fused" in tkey) else 1.0 / (n_blstm * T))
        roof = {"bound": "tensor", "kernel": kname,
                "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak, "traffic": traffic,
                "peak_source": pk["src"] + " (sustained bf16)",
                "us_per_launch": k_ms * 1e3 / n_launch_k, "launches": n_launch_k,
                "algorithmic_flops_per_launch": k_flops * n * steps / n_launch_k,
                "tensor_pipe_frac_with_3_passes": 3 * achieved / peak,
                "stage_ms_per_step": {k: v / steps for k, v in stage4.items()},
                "stage_tflops": {k: stages[k].get("achieved_tflops") for k in stage4},
                "whole_step": {"algorithmic_mflop_per_window": sum(flops.values()) / 1e6,
                               "achieved_tflops": sum(flops.values()) * n / (ms_step * 1e-3) / 1e12,
                               "frac": sum(flops.values()) * n / (ms_step * 1e-3) / 1e12 / peak},
                "hbm_peak_gbs": pk["hbm_gbs"]}
    return roof, stages


def bench_windows(cx, args, steps, warmup, with_e2e=True, with_cpu=True):
    """Weak-scaling replica line: every rank runs its own ``--windows`` pre-windowed synthetic windows (uniform
    1000-window reads); no data crosses ranks."""
    import torch
    from nanopore_rnn_b200 import engine
    rank, world, local, dev, dist, m, net = cx.rank, cx.world, cx.local, cx.dev, cx.dist, cx.m, cx.net
    barrier, max_over_ranks, sum_over_ranks = cx.barrier, cx.max_over_ranks, cx.sum_over_ranks
    args = argparse.Namespace(**dict(vars(args), steps=steps, warmup=warmup, no_e2e=not with_e2e, no_cpu=not with_cpu))

    # each rank owns its own reads (weak scaling): no collective on the data path
    x, lens, goff = make_workload(args.windows, seed=100 + rank)
    n = len(x)
    samples = int(lens.sum())
    ctc = WL["graph"] == "CtcLoss"

    # ---- device-resident job: inputs already in HBM when the timed region starts
    x_pin = torch.from_numpy(x).pin_memory()
    lens_pin = torch.from_numpy(lens).pin_memory()
    x_d, lens_d = x_pin.to(dev), lens_pin.to(dev)
    cap = args.chunk_windows or m.max_chunk_windows()
    if goff is None:        # no BN groups (pure BLSTM network): any cut is legal -> equal chunks under the cap
        cuts = np.unique(np.linspace(0, n, -(-n // cap) + 1).astype(np.int64))
        plan_off, chunks = cuts, [(i, i + 1) for i in range(len(cuts) - 1)]
        chunk_goff = [None] * len(chunks)
    else:
        plan_off = goff
        chunks = engine.plan_chunks(goff, cap, m.chunk_align())
        chunk_goff = [torch.from_numpy(np.ascontiguousarray(goff[g:h + 1] - goff[g], dtype=np.int32)).to(dev)
                      for g, h in chunks]
    labels_d = torch.empty(n * T, dtype=torch.int32, device=dev)
    llen_d = torch.empty(n, dtype=torch.int32, device=dev) if ctc else None
    launches = [0]

    def step():
        for (g, h), go in zip(chunks, chunk_goff):
            w0, w1 = int(plan_off[g]), int(plan_off[h])
            m.forward_device(x_d[w0:w1], lens_d[w0:w1], go, labels_out=labels_d[w0 * T:w1 * T],
                             label_len_out=llen_d[w0:w1] if ctc else None)
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
               "h2d_bytes_per_step": int(sum_over_ranks(x.nbytes + lens.nbytes + (goff.nbytes if goff is not None else 0))),
               "d2h_bytes_per_step": int(sum_over_ranks(labels.nbytes + (llen.nbytes if llen is not None else 0))),
               "ms_per_step": dt * 1e3}

    roof, stages = roofline_of(cx, stage, n, len(chunks), max(args.steps, 1), ms_step)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_baseline(args.cpu_windows)
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "windows_per_gpu": n, "seq_len": T, "bn_group": BATCH,
                       "n_classes": NCLASS, "chunks_per_step": len(chunks),
                       "l2": "inputs (%.0f MB) and activations larger than L2" % (x.nbytes / 1e6),
                       "parallelism": "independent replicas: every rank runs its own windows on %d GPU(s), no collective" % world},
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(n_launch), "roofline": roof,
            "roofline_stages": stages, "cpu_baseline": cpu}
    del x_d, lens_d, labels_d, llen_d
    return line


# ---------------------------------------------------------------------------- the sharded read set (configs[1] + [3])
READSET = {"n_reads": 65536, "seed": 20, "mu_samples": 40000.0, "sigma": 0.8, "clip": (2000, 400000)}


def make_read_set(total_windows, step):
    """65 536 read lengths ~ LogNormal(ln 40 000, 0.8) clipped to [2 000, 400 000] samples
    (seed 20).  One bench step can not take all of them (36 M windows at step 100 = a minute per step on one GPU), so
    the job is the first reads of that set whose windows add up to ``total_windows`` (the last one cut to fit):
    the 1 048 576-window job of configs[1], made of config 4's variable-length reads."""
    from nanopore_rnn_b200 import synthetic as syn
    from nanopore_rnn_b200.job import window_counts
    lengths = syn.lognormal_read_lengths(READSET["n_reads"], READSET["seed"], np.log(READSET["mu_samples"]),
                                         READSET["sigma"], *READSET["clip"])
    cum = np.cumsum(window_counts(lengths, step))
    k = int(np.searchsorted(cum, total_windows, side="left")) + 1
    lengths = lengths[:k].copy()
    over = int(cum[k - 1] - total_windows) if k <= len(cum) else 0
    if over > 0:
        lengths[-1] -= over * step
    return lengths


def synth_read(read_id, length):
    """int16 current samples of one synthetic read: N(500, 60) around a slow drift; seeded by the read id alone, so a
    read's samples do not depend on how the set is sharded."""
    rng = np.random.default_rng([7, int(read_id)])
    v = rng.standard_normal(int(length), dtype=np.float32) * 60.0 + 500.0 + 20.0 * np.sin(np.arange(int(length)) * 1e-3)
    return np.rint(v).astype(np.int16)


def bench_reads(cx, args):
    """Strong scaling: ONE read set, reads assigned to ranks by LPT on window counts (shard.assign_reads), each rank
    packs its share length-sorted and runs the device-resident chain (job.ReadJob); the end-to-end leg gathers the
    called reads to rank 0 in original read order inside the timed region (job.gather_consensus)."""
    import torch
    from nanopore_rnn_b200 import shard
    from nanopore_rnn_b200.job import ReadJob, gather_consensus, window_counts
    rank, world, dev, m = cx.rank, cx.world, cx.dev, cx.m
    lengths = make_read_set(args.windows, args.step)
    counts = window_counts(lengths, args.step)
    assignment = shard.assign_reads(counts, world)
    mine = np.asarray(assignment[rank], dtype=np.int64)
    job = ReadJob(m, step=args.step, start_index=0, batch_size=BATCH, chunk_windows=args.chunk_windows or None)
    plan, host = job.stage_raw([synth_read(r, lengths[r]) for r in mine])
    read_ids = mine[plan.order]                                  # packed position -> original read index
    samples = plan.valid_samples
    total_samples = cx.sum_over_ranks(samples)
    raw_d = torch.from_numpy(host).to(dev)
    steps = max(args.steps, 1)

    # ---- device-resident: samples already in HBM; pack -> forward -> assemble, CUDA events on the chain's stream
    for _ in range(args.warmup):
        job.run_device(raw_d, plan)
    cx.barrier()
    sampler = ClockSampler(cx.local)
    sampler.start()
    m.set_profile(True)
    n_launch = 0
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(steps):
        job.run_device(raw_d, plan)
        n_launch += job.last["launches"]
    ev1.record()
    cx.barrier()
    clocks = sampler.finish()
    my_ms = ev0.elapsed_time(ev1) / steps
    ms_step = cx.max_over_ranks(my_ms)
    per_rank_ms = cx.all_values(my_ms)
    stage = m.stage_ms()
    m.set_profile(False)
    value = total_samples / (ms_step * 1e-3)

    # ---- end to end: pinned host int16 samples in, called reads (FASTA bodies) of ALL ranks out on rank 0
    e2e = None
    if not args.no_e2e:
        raw_e = torch.empty_like(raw_d)
        host_t = torch.from_numpy(host)
        gather_s = [0.0]
        called = [None]

        enq_s = [0.0]

        gpu_ms = [0.0]
        gpu_steps = []
        phases = []
        ge0, ge1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

        def e2e_step(timed):
            t_a = time.perf_counter()
            ge0.record()
            raw_e.copy_(host_t, non_blocking=True)
            cons, cons_off, status = job.run_device(raw_e, plan)
            ge1.record()
            if timed:
                t_b = time.perf_counter()
                enq_s[0] += t_b - t_a                          # host time to enqueue the chain (the assembly's table upload waits for the forward)
                if e2e_mode == "step":
                    time.sleep(min(0.3, 0.4e-3 * my_ms))       # (well inside this rank's step: the GPU is under load when sampled)
                    sampler_e.sample_once()
                torch.cuda.current_stream(dev).synchronize()
                t = time.perf_counter()
                gpu_ms[0] += ge0.elapsed_time(ge1)             # device time of copy + chain inside this step
                gpu_steps.append(round(ge0.elapsed_time(ge1), 1))
            called[0] = gather_consensus(cons, cons_off, plan.n_reads, read_ids, len(lengths), rank, world, dev)
            if timed:
                t_c = time.perf_counter()
                gather_s[0] += t_c - t
                phases.append((round((t_b - t_a) * 1e3, 1), round((t - t_b) * 1e3, 1), round((t_c - t) * 1e3, 1)))

        for _ in range(2):
            e2e_step(False)
        cx.barrier()
        # No NVML polling in this leg: its queries were measured to take up to a second now and then while the GPU is under
        # load, and -- from a polling thread as well as from this one -- to delay the host's wake-up from the stream
        # synchronisation by 0.1-0.9 s (profiles/r02_e2e_nvml_hiccups.txt, tools/hic.sh).  The device-resident leg above is
        # timed with CUDA events and keeps its sampler; this leg reports the device time of its own steps instead
        # (NTB_E2E_SAMPLER=thread|step re-enables the sampling for A/B runs).
        sampler_e = ClockSampler(cx.local)
        e2e_mode = os.environ.get("NTB_E2E_SAMPLER", "none")
        e2e_thread = e2e_mode == "thread"
        if e2e_thread:
            sampler_e.start()
        t0 = time.perf_counter()
        step_s = []
        for _ in range(steps):
            t_s = time.perf_counter()
            e2e_step(True)
            step_s.append(time.perf_counter() - t_s)
        torch.cuda.synchronize()
        t1 = time.perf_counter()                 # this rank's K steps are done; the slowest rank's elapsed time is the job's
        cx.barrier()
        t2 = time.perf_counter()
        dt = cx.max_over_ranks(t1 - t0) / steps
        clocks_e = sampler_e.finish()
        if os.environ.get("NTB_BENCH_DEBUG"):
            sys.stderr.write("e2e step wall ms: %s device ms: %s (enqueue, wait, gather) ms: %s; loop %.1f ms, closing barrier %.1f ms\n" % ([round(v * 1e3, 1) for v in step_s], gpu_steps, phases, (t1 - t0) * 1e3, (t2 - t1) * 1e3))
        d2h = sum(len(c) for c in called[0]) + 16 * len(lengths) + 8 if rank == 0 else 0
        e2e = {"value": total_samples / dt, "unit": UNIT, "ms_per_step": dt * 1e3,
               "h2d_bytes_per_step": int(cx.sum_over_ranks(host.nbytes + 16 * (plan.n_reads + 1) + 4 * len(plan.group_off))),
               "d2h_bytes_per_step": int(cx.sum_over_ranks(d2h)),
               # time inside gather_consensus: on rank 0 (sizes exchange + NCCL gather + one D2H + host reorder; includes
               # waiting for the slowest rank to arrive) and the smallest over ranks (the exchange itself, no waiting)
               "gather_ms_per_step": cx.all_values(gather_s[0])[0] / steps * 1e3,
               "gather_ms_per_step_min_rank": min(cx.all_values(gather_s[0])) / steps * 1e3,
               "enqueue_ms_per_step": cx.max_over_ranks(enq_s[0]) / steps * 1e3,
               "sm_mhz": clocks_e.get("sm_mhz"), "reasons": clocks_e.get("reasons"),
               "device_ms_per_step": cx.max_over_ranks(gpu_ms[0]) / steps,
               "called_bases_per_step": int(sum(len(c) for c in called[0])) if rank == 0 else None,
               "what": "pinned int16 samples -> H2D -> z-score/windows -> forward -> consensus assembly -> NCCL gather "
                       "of the called reads to rank 0 -> one D2H -> original read order"}
        if rank == 0:
            assert all(c is not None for c in called[0]), "a read is missing from the gather"

    roof, stages = roofline_of(cx, stage, plan.n_windows, len(plan.chunks), steps, my_ms)
    win_rank = [int(v) for v in cx.all_values(float(plan.n_windows))]
    reads_rank = [int(v) for v in cx.all_values(float(plan.n_reads))]
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_baseline(args.cpu_windows)
        cpu["sample"] += "; forward only (the CPU leg does not pay for z-score/windowing or the consensus assembly)"
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "%s, as ONE read set of BASELINE configs[3]: first %d reads of the 65 536-read "
                                   "LogNormal(ln 40000, 0.8) set (SURVEY 8d config 4), %d windows in total, window step %d, "
                                   "reads LPT-sharded over %d GPU(s), length-sorted packing"
                                   % (WORKLOAD, len(lengths), int(counts.sum()), args.step, world),
                       "reads": len(lengths), "total_windows": int(counts.sum()), "total_samples": int(lengths.sum()),
                       "seq_len": T, "step": args.step, "bn_group": BATCH, "n_classes": NCLASS,
                       "windows_per_rank": win_rank, "reads_per_rank": reads_rank,
                       "imbalance_max_over_mean": max(win_rank) / (sum(win_rank) / len(win_rank)),
                       "device_ms_per_rank": per_rank_ms, "chunks_per_step": len(plan.chunks),
                       "l2": "inputs (%.0f MB int16, %.0f MB of windows) and activations larger than L2"
                             % (host.nbytes / 1e6, plan.n_windows * T * 4 / 1e6),
                       "parallelism": "reads sharded over %d GPU(s) by LPT on window counts; no collective on the compute "
                                      "path, one NCCL gather of the called reads (e2e)" % world},
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(n_launch), "roofline": roof,
            "roofline_stages": stages, "cpu_baseline": cpu}
    del raw_d
    job._bufs.clear()
    return line


def run_b200(args):
    cx = Ctx()
    reads_ok = WL["graph"] == "CtcLoss" and WL["len_x"] == 1
    if args.workload == "reads" and reads_ok:
        line = bench_reads(cx, args)
        if not args.no_replica:
            # second key: the weak-scaling replica line of round 1 (uniform pre-windowed windows, every rank its own),
            # device-resident only, over min(steps, 3) steps after 1 warm-up -- context for the strong-scaling value
            cx.torch.cuda.empty_cache()
            rl = bench_windows(cx, args, steps=min(max(args.steps, 1), 3), warmup=1, with_e2e=False, with_cpu=False)
            line["replica_weak"] = {"value": rl["value"], "unit": UNIT, "ms_per_step": rl["ms_per_step"], "scaling": "weak",
                                    "steps": rl["steps"], "warmup": rl["warmup"], "windows_per_gpu": rl["config"]["windows_per_gpu"],
                                    "what": rl["config"]["parallelism"]}
    else:
        line = bench_windows(cx, args, args.steps, args.warmup, with_e2e=not args.no_e2e, with_cpu=not args.no_cpu)
    if cx.rank == 0:
        print(json.dumps(line))
    cx.close()
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
