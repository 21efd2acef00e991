"""Diagnostic (not a test): logits error of the GPU path vs the fp32 and fp64 oracle as BLSTM layers stack."""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_np as A
import nanopore_rnn_b200  # noqa
from nanopore_rnn_b200 import engine, synthetic as syn

H = int(sys.argv[1]) if len(sys.argv) > 1 else 256
T, Cc, n = 100, 1025, 24
rng = np.random.default_rng(30)
x = rng.standard_normal((n, T, 12)).astype(np.float32)
lens = rng.integers(1, T + 1, n).astype(np.int32)
for L in (1, 2, 3, 5):
    net = [{"bias": 5.0, "type": "blstm", "name": "blstm_layer%d" % i, "size": H, "concat": True} for i in range(1, L + 1)]
    w = syn.random_weights(net, 12, Cc, seed=31, graph="CrossEntropy")
    m = engine.Model(net, w, "CrossEntropy", T, 12, Cc, device="cuda:0")
    labels, llen, logits = m.forward_host(x, lens, None, want_logits=True)
    ref32, _ = A.forward_batch(x, lens, net, w, "CrossEntropy")
    try:
        ref64, _ = A.forward_batch(x, lens, net, w, "CrossEntropy", dtype=np.float64)
    except TypeError:
        ref64 = None
    e32 = np.abs(logits - ref32.reshape(-1, Cc)).max()
    msg = "H=%d layers=%d  |gpu-ref32|=%.3e" % (H, L, e32)
    if ref64 is not None:
        msg += "  |gpu-ref64|=%.3e  |ref32-ref64|=%.3e" % (np.abs(logits - ref64.reshape(-1, Cc)).max(),
                                                         np.abs(ref32 - ref64).max())
    print(msg, flush=True)
