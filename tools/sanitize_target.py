#!/usr/bin/env python
"""The small job `tools/sanitize.sh` runs under compute-sanitizer (memcheck / racecheck / synccheck): one forward of
each tensor-core path through the C ABI -- the CG=2 persistent GEMM (`k_gemm_tc<2>`), the fused-transform conv
(`k_conv_fused`), the TMEM-resident recurrence (`k_lstm_recurrence_tc`), the per-timestep H=256 recurrence -- and one
call of each widened row (event features, event detection, .signal parse, consensus assembly).  Results are checked
against the committed goldens so that a sanitizer run that "passes" on garbage cannot slip through."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import nanopore_rnn_b200  # noqa: E402,F401
from nanopore_rnn_b200 import assembly, dataset, engine, eventdetection, features, synthetic as syn  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "all"
gd = os.path.join(ROOT, "tests", "golden")

if which in ("all", "forward"):
    g = np.load(os.path.join(gd, "crnn_run_nanotensor.npz"))
    net, Cn = syn.NETWORKS["run_nanotensor"], int(g["n_classes"])
    w = syn.random_weights(net, 1, Cn, seed=int(g["seed"]) + 2)
    m = engine.Model(net, w, "CtcLoss", 100, 1, Cn, device="cuda:0")
    labels, llen, logits = m.forward_host(g["x"], g["lens"], g["group_off"], want_logits=True)
    err = float(np.abs(logits - g["logits"].reshape(-1, Cn)).max())
    print("forward run_nanotensor golden: %d windows, %d launches, max|logit err| %.3g" % (len(g["x"]), m.last_launches(), err))
    assert err <= 1e-3
    # a job large enough for several persistent tiles per CTA pair and a ragged tail tile
    x, lens, _ = syn.synthetic_reads(3, 130, 100, seed=5)
    goff = engine.group_offsets([130] * 3, 10)
    a = m.forward_host(x, lens, goff, want_logits=True)[2]
    b = m.forward_host(x, lens, goff, want_logits=True)[2]
    assert np.array_equal(a, b)
    print("forward 390 windows: bit-stable run to run")
    m.close()

if which in ("all", "stress"):
    net = syn.NETWORKS["stress_blstm256"][:2]
    w = syn.random_weights(net, 12, 64, seed=31, graph="CrossEntropy")
    rng = np.random.default_rng(3)
    x = rng.standard_normal((40, 20, 12)).astype(np.float32)
    lens = rng.integers(0, 21, 40).astype(np.int32)
    m = engine.Model(net, w, "CrossEntropy", 20, 12, 64, device="cuda:0")
    a = m.forward_host(x, lens, None, want_logits=True)[2]
    print("forward 2 x BLSTM(256) T=20: %d launches, finite %s" % (m.last_launches(), bool(np.isfinite(a).all())))
    m.close()

if which in ("all", "rows"):
    ge = np.load(os.path.join(gd, "event_features.npz"))
    off = ge["read_off"]
    reads = [{k: ge[k][a:b] for k in ("mean", "stdv", "length")} for a, b in zip(off[:-1], off[1:])]
    X, _ = features.events_to_features_batch(reads)
    assert np.array_equal(X, ge["features"])
    print("event features: bit-identical to the reference golden")
    gd_ = np.load(os.path.join(gd, "event_detect.npz"))
    o, eo = gd_["read_off"], gd_["event_off_0"]
    ev = eventdetection.minknow_event_detect(gd_["raw"][o[0]:o[1]], float(gd_["sample_rate"]), [3, 6], [1.4, 1.1], 0.2)
    want = gd_["events_0"][eo[0]:eo[1]]
    assert len(ev) == len(want) and np.array_equal(ev["mean"], want[:, 2])
    print("event detection: %d events identical" % len(ev))
    ga = np.load(os.path.join(gd, "assembly.npz"))
    seqs = assembly.assemble_reads(ga["labels"], ga["read_off"])
    assert "".join(seqs) == ga["consensus"].tobytes().decode()
    print("assembly: %d reads identical" % len(seqs))
    import tempfile
    rng = np.random.default_rng(0)
    with tempfile.TemporaryDirectory() as td:
        paths = []
        for i, n in enumerate([5000, 1, 12345]):
            p = os.path.join(td, "r%d.signal" % i)
            vals = rng.integers(-300, 900, n)
            open(p, "w").write(" ".join(str(int(v)) for v in vals))
            paths.append((p, vals))
        s, ro = dataset.read_signals_gpu([p for p, _ in paths])
        assert np.array_equal(s, np.concatenate([v for _, v in paths]).astype(np.int16))
    print("signal parse: %d samples identical" % len(s))
print("sanitize target OK")
