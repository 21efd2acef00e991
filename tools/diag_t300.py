"""Diagnostic (not a test): where the logits error of the Chiron shape (3 residual + 3 BLSTM(100) + chiron_fnn) comes from
as T and the layer counts grow -- GPU and the C++ float32 oracle, both against the NumPy restatement evaluated in float64."""
import os
import sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_np as A
import oracle_cpu as Cc
import nanopore_rnn_b200  # noqa
from nanopore_rnn_b200 import engine, synthetic as syn

Cn = 5
for T in (100, 300):
    for nres, nb in ((3, 3), (3, 1), (1, 1), (1, 3)):
        net = [{"type": "residual_layer", "name": "res%d" % i, "out_channel": 256, "batchnorm": i == 1} for i in range(1, nres + 1)] + \
              [{"type": "blstm", "name": "blstm%d" % i, "size": 100, "bias": 5.0, "concat": True} for i in range(1, nb + 1)] + \
              [{"type": "chiron_fnn", "name": "fnn", "size": 64, "bias": 1.0}]
        w = syn.random_weights(net, 1, Cn, seed=301)
        groups = [30, 30, 11]
        n = sum(groups)
        x, lens, _ = syn.synthetic_reads(1, n, T, seed=302)
        goff = np.concatenate([[0], np.cumsum(groups)]).astype(np.int32)
        m = engine.Model(net, w, "CtcLoss", T, 1, Cn, device="cuda:0")
        labels, llen, logits = m.forward_host(x, lens, goff, want_logits=True)
        ref, _, _ = Cc.forward(x, lens, net, w, "CtcLoss", group_off=goff)
        ref64 = np.concatenate([A.forward_batch(x[a:b], lens[a:b], net, w, "CtcLoss", dtype=np.float64)[0].reshape(-1, Cn)
                                for a, b in zip(goff[:-1], goff[1:])])
        d = np.abs(logits - ref64)
        print("T=%d res=%d blstm=%d  |gpu-f64| max %.3e mean %.3e   |cpp32-f64| max %.3e   logits scale %.2f"
              % (T, nres, nb, d.max(), d.mean(), np.abs(ref.reshape(ref64.shape) - ref64).max(), np.abs(ref64).max()), flush=True)
        m.close()
