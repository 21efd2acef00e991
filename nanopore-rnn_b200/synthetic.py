"""Synthetic workloads of the shapes SURVEY.md section 8(d) names: random-init weights in the
TF checkpoint layouts and synthetic reads / windows.  Used by bench.py, smoke() and the tests
(there is no network for datasets or trained checkpoints)."""
import numpy as np


def random_weights(network, len_x, n_classes, seed, graph="CtcLoss", readout_gain=1.0):
    """Random-init weights with the shapes/initialiser families the reference creates
    (LSTM kernel glorot-uniform over [D+H,4H], bias zeros -- TF 1.4 LSTMCell defaults;
    conv xavier-uniform network.py:316; BN scale/offset network.py:344-347 -- drawn here as
    N(1,.1)/N(0,.1) so BN is non-degenerate; chiron_fnn network.py:167-171; dense :298-301).
    LSTM biases get small noise instead of exact zeros so bias handling is exercised."""
    rng = np.random.default_rng(seed)
    f32 = np.float32
    layers = []
    D = len_x
    for layer in network:
        t = layer["type"]
        if t == "blstm":
            H = layer["size"]
            lim = np.sqrt(6.0 / (D + H + 4 * H))
            w = {}
            for d in ("fw", "bw"):
                w["kernel_" + d] = rng.uniform(-lim, lim, (D + H, 4 * H)).astype(f32)
                w["bias_" + d] = (0.05 * rng.standard_normal(4 * H)).astype(f32)
            layers.append(w)
            D = 2 * H
        elif t == "residual_layer":
            C = layer["out_channel"]

            def xav(k, cin, cout):
                lim = np.sqrt(6.0 / (k * cin + k * cout))
                return rng.uniform(-lim, lim, (1, k, cin, cout)).astype(f32)

            w = {"conv1": xav(1, D, C), "conv2a": xav(1, D, C), "conv2b": xav(3, C, C), "conv2c": xav(1, C, C)}
            names = ["conv2a", "conv2b", "conv2c"] + (["conv1"] if layer.get("batchnorm") else [])
            for c in names:
                w[c + "_bn_scale"] = (1.0 + 0.1 * rng.standard_normal(C)).astype(f32)
                w[c + "_bn_offset"] = (0.1 * rng.standard_normal(C)).astype(f32)
            layers.append(w)
            D = C
        elif t == "chiron_fnn":
            H = D // 2
            layers.append({"weights": (rng.standard_normal((2, H)) * np.sqrt(2.0 / D)).astype(f32),
                           "bias": (0.05 * rng.standard_normal(H)).astype(f32)})
            D = H
        else:
            raise ValueError(t)
    # readout_gain > 1 spreads the logits so that the argmax labels are diverse (an untrained
    # readout is nearly constant, which would make label-agreement tests vacuous)
    out_w = (readout_gain * rng.standard_normal((D, n_classes)) * np.sqrt(2.0 / (2 * n_classes))).astype(f32)
    out_b = (0.05 * rng.standard_normal(n_classes)).astype(f32)
    return {"layers": layers, "out_w": out_w, "out_b": out_b}


def synthetic_reads(n_reads, windows_per_read, seq_len, seed, len_x=1, ragged_last=True):
    """Windows of ``n_reads`` z-scored synthetic reads: i.i.d. N(0,1)
    per read then z-scored over the read; every window has ``seq_len`` valid samples except the
    last window of each read, whose length ~ U{1..seq_len} (zero padded).

    Returns x float32 [n_reads*windows_per_read, seq_len(, len_x)], lens int32, read_id int32."""
    rng = np.random.default_rng(seed)
    n = n_reads * windows_per_read
    shape = (n_reads, windows_per_read * seq_len) + ((len_x,) if len_x > 1 else ())
    sig = rng.standard_normal(shape, dtype=np.float32)
    lens = np.full((n_reads, windows_per_read), seq_len, dtype=np.int32)
    if ragged_last:
        lens[:, -1] = rng.integers(1, seq_len + 1, size=n_reads)
    if len_x == 1:
        valid = (windows_per_read - 1) * seq_len + lens[:, -1]
        mask = np.arange(shape[1])[None, :] < valid[:, None]
        cnt = mask.sum(1, keepdims=True)
        mean = (sig * mask).sum(1, keepdims=True) / cnt
        var = (((sig - mean) * mask) ** 2).sum(1, keepdims=True) / cnt
        sig = ((sig - mean) / np.sqrt(var)) * mask
        x = sig.reshape(n, seq_len).astype(np.float32)
    else:
        x = sig.reshape(n, seq_len, len_x)
        m = (np.arange(seq_len)[None, :] < lens.reshape(-1)[:, None])
        x = x * m[:, :, None]
    read_id = np.repeat(np.arange(n_reads, dtype=np.int32), windows_per_read)
    return x, lens.reshape(-1), read_id


def lognormal_read_lengths(n_reads, seed, mu=np.log(40000.0), sigma=0.8, lo=2000, hi=400000):
    """Read lengths of SURVEY.md 8(d) config 4."""
    rng = np.random.default_rng(seed)
    return np.clip(rng.lognormal(mu, sigma, n_reads), lo, hi).astype(np.int64)


NETWORKS = {
    # configs/run_nanotensor.config.json:29-76 of the reference
    "run_nanotensor": [
        {"type": "residual_layer", "name": "residual_layer_1", "out_channel": 256, "batchnorm": True},
        {"type": "residual_layer", "name": "residual_layer_2", "out_channel": 256, "batchnorm": False},
        {"type": "residual_layer", "name": "residual_layer_3", "out_channel": 256, "batchnorm": False},
        {"bias": 5.0, "type": "blstm", "name": "blstm_layer1", "size": 100, "concat": True},
        {"bias": 5.0, "type": "blstm", "name": "blstm_layer2", "size": 100, "concat": True},
        {"bias": 5.0, "type": "blstm", "name": "blstm_layer3", "size": 100, "concat": True},
        {"bias": 1.0, "type": "chiron_fnn", "name": "chiron_fnn_layer", "size": 64}],
    # configs/signal_CRNN.config.json:24-51
    "signal_CRNN": [
        {"type": "residual_layer", "name": "residual_layer_1", "out_channel": 256, "batchnorm": True},
        {"type": "residual_layer", "name": "residual_layer_2", "out_channel": 256, "batchnorm": False},
        {"bias": 5.0, "type": "blstm", "name": "blstm_layer1", "size": 100, "concat": True},
        {"bias": 1.0, "type": "chiron_fnn", "name": "chiron_fnn_layer", "size": 64}],
    # nanotensor/tests/run_nanotensor_test.py:42 (the net the shipped test checkpoint was trained with)
    "event_fixture": [
        {"bias": 5.0, "type": "blstm", "name": "blstm_layer1", "size": 5, "concat": True}],
    # BASELINE.json configs[4]
    "stress_blstm256": [
        {"bias": 5.0, "type": "blstm", "name": "blstm_layer%d" % i, "size": 256, "concat": True}
        for i in range(1, 6)],
}
