"""TEST INFRASTRUCTURE -- run-time probe for the REAL reference arithmetic (TensorFlow).

The reference's forward is a chain of TensorFlow 1.4 ops (nanotensor/network.py); TF is not in this image, so the
CPU oracles under oracle/ are restatements and forward parity is "unpinned" (DESIGN.md section 3).  BASELINE.md /
SURVEY.md section 7 promise a probe at run time: ``probe()`` tries ``import tensorflow`` where the code actually runs
(bench.py, smoke(), the tests -- here AND on the GPU box) and reports what it OBSERVED instead of a hard-coded string.

If TensorFlow ever imports, ``pin_against_goldens(golden_dir)`` builds the reference's graph out of the same TF ops the
reference calls (``tf.compat.v1`` spelling of network.py:266-349, 507-533: ``rnn_cell.LSTMCell``,
``bidirectional_dynamic_rnn``, ``nn.conv2d``, ``nn.moments``, ``nn.batch_normalization``) on the committed golden
inputs + weights and returns the max |logit| distance to the committed goldens: that number PINS the restatements
(<= 1e-4 expected).  It has never run (TF absent everywhere this repo has been); it is exercised by
tests/test_oracle.py only when the import succeeds.

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs import this module.
"""
import os

import numpy as np


def probe():
    """-> {"importable": bool, "version": str|None, "error": str|None}; never raises."""
    try:
        import tensorflow as tf  # noqa: F401
        return {"importable": True, "version": getattr(tf, "__version__", "?"), "error": None}
    except Exception as e:  # ImportError, or a broken install
        return {"importable": False, "version": None, "error": "%s: %s" % (type(e).__name__, str(e)[:120])}


def describe():
    p = probe()
    if p["importable"]:
        return "tensorflow %s importable on this box" % p["version"]
    return "TensorFlow not importable on this box (observed: %s)" % p["error"]


# ------------------------------------------------------------------------------------------------------------------
# The reference graph out of TF ops (only reachable when TensorFlow imports)
# ------------------------------------------------------------------------------------------------------------------
def _tf_forward(x, lens, network, weights, graph):
    import tensorflow as tf
    v1 = tf.compat.v1
    v1.disable_eager_execution()
    g = tf.Graph()
    with g.as_default():
        xs = v1.placeholder(tf.float32, shape=x.shape)
        ls = v1.placeholder(tf.int32, shape=lens.shape)
        v = tf.reshape(xs, [x.shape[0], x.shape[1], 1]) if graph == "CtcLoss" else xs          # network.py:515-518
        T = x.shape[1]
        assign = []

        def bn(y, scale, offset):                                                             # network.py:339-349
            mean, var = tf.nn.moments(y, [0, 1, 2])
            return tf.nn.batch_normalization(y, mean, var, tf.constant(offset), tf.constant(scale), 1e-5)

        def conv(y, W):                                                                       # network.py:312-329
            return tf.nn.conv2d(y, tf.constant(W), strides=[1, 1, 1, 1], padding="SAME")

        for li, (layer, w) in enumerate(zip(network, weights["layers"])):
            if layer["type"] == "residual_layer":                                             # network.py:385-400
                cin = int(v.shape[2])
                y = tf.reshape(v, [-1, 1, T, cin])
                a = conv(y, w["conv1"])
                if layer.get("batchnorm"):
                    a = bn(a, w["conv1_bn_scale"], w["conv1_bn_offset"])
                b = tf.nn.relu(bn(conv(y, w["conv2a"]), w["conv2a_bn_scale"], w["conv2a_bn_offset"]))
                b = tf.nn.relu(bn(conv(b, w["conv2b"]), w["conv2b_bn_scale"], w["conv2b_bn_offset"]))
                b = bn(conv(b, w["conv2c"]), w["conv2c_bn_scale"], w["conv2c_bn_offset"])
                v = tf.reshape(tf.nn.relu(a + b), [-1, T, int(w["conv2c"].shape[-1])])
            elif layer["type"] == "blstm":                                                    # network.py:266-288
                H = int(layer["size"])
                with v1.variable_scope("blstm_%d" % li):
                    fw = v1.nn.rnn_cell.LSTMCell(H, forget_bias=float(layer["bias"]), state_is_tuple=True)
                    bw = v1.nn.rnn_cell.LSTMCell(H, forget_bias=float(layer["bias"]), state_is_tuple=True)
                    outs, _ = v1.nn.bidirectional_dynamic_rnn(fw, bw, v, dtype=tf.float32, sequence_length=ls)
                    v = tf.concat(outs, 2)
                for cell, d in ((fw, "fw"), (bw, "bw")):
                    k, b = cell.variables[0], cell.variables[1]
                    assign += [v1.assign(k, w["kernel_" + d]), v1.assign(b, w["bias_" + d])]
            elif layer["type"] == "chiron_fnn":                                               # network.py:164-175
                H = int(v.shape[2]) // 2
                r = tf.reshape(v, [-1, T, 2, H])
                v = tf.reduce_sum(r * tf.constant(w["weights"]), axis=2) + tf.constant(w["bias"])
            else:
                raise ValueError(layer["type"])
        F = int(v.shape[2])
        logits = tf.matmul(tf.reshape(v, [-1, F]), tf.constant(weights["out_w"])) + tf.constant(weights["out_b"])
        with v1.Session(graph=g) as sess:
            sess.run(v1.global_variables_initializer())
            sess.run(assign)
            return sess.run(logits, {xs: x, ls: lens})


def pin_against_goldens(golden_dir):
    """Max |TF logits - committed golden logits| per fixture; raises if TensorFlow is not importable."""
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    root = os.path.dirname(here)
    for p in (root, here):
        if p not in sys.path:
            sys.path.insert(0, p)
    import nanopore_rnn_b200  # noqa: F401
    from nanopore_rnn_b200 import checkpoint as ckpt, synthetic as syn
    out = {}
    g = np.load(os.path.join(golden_dir, "event_fixture.npz"))
    net = syn.NETWORKS["event_fixture"]
    w = ckpt.weights_for_network(ckpt.load_checkpoint(os.path.join(golden_dir, "test_model-30")), net)
    lg = _tf_forward(g["x"], g["lens"], net, w, "CrossEntropy")
    out["event_fixture"] = float(np.abs(lg[:240] - g["logits_head"]).max())
    for name, key in (("crnn_signal", "signal_CRNN"), ("crnn_run_nanotensor", "run_nanotensor")):
        g = np.load(os.path.join(golden_dir, name + ".npz"))
        net, Cn = syn.NETWORKS[key], int(g["n_classes"])
        w = syn.random_weights(net, 1, Cn, seed=int(g["seed"]) + 2)
        go = g["group_off"]
        lg = np.concatenate([_tf_forward(g["x"][a:b], g["lens"][a:b], net, w, "CtcLoss")
                             for a, b in zip(go[:-1], go[1:])])
        out[name] = float(np.abs(lg - g["logits"].reshape(-1, Cn)).max())
    return out
