"""CPU ORACLE (numpy) -- TEST INFRASTRUCTURE ONLY.

A restatement of the reference's TensorFlow-1.4 forward for the NanoTensor BLSTM basecaller
(nanotensor/network.py as driven by nanotensor/run_nanotensor.py).  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` leg may import
this; the product path (``nanopore-rnn_b200/``) never does.

PARITY UNPINNED: the reference ships no golden logits/labels for this path (both hot-path test
files are fully commented out: nanotensor/tests/network_test.py:14-110,
nanotensor/tests/run_nanotensor_test.py:13-66) and its arithmetic lives in tensorflow==1.4.0
(requirements35.yml:80), absent here.  This file restates TF 1.4's published op semantics at the
reference's call sites and is cross-checked against an independent torch build
(oracle/oracle_torch.py); the fixtures under tests/golden/ are THIS oracle's outputs on the
reference's shipped inputs + weights (scripts: tests/golden/make_golden.py).

All arithmetic is float32 (network.py:281 ``dtype=tf.float32``) unless ``dtype=np.float64`` is
passed for error budgeting.  Labels are int64 (tf.argmax / ctc decoder values).
"""
import difflib

import numpy as np


# ------------------------------------------------------------------ elementary maps
def sigmoid(x):
    # tf.sigmoid == 1/(1+exp(-x)); evaluated in the array's own dtype
    one = x.dtype.type(1)
    return one / (one + np.exp(-x))


# ------------------------------------------------------------------ a3: blstm (network.py:266-288)
def lstm_dir(v, lens, kernel, bias, forget_bias, reverse):
    """One direction of tf.nn.bidirectional_dynamic_rnn over rnn.LSTMCell (network.py:275-282).

    TF 1.4 LSTMCell.call (no peepholes / projection / clipping):
        z = [x_t, h_{t-1}] @ kernel + bias ; i, j, f, o = split(z, 4)
        c = sigmoid(f + forget_bias) * c_prev + sigmoid(i) * tanh(j) ; h = sigmoid(o) * tanh(c)
    dynamic_rnn(sequence_length): for t >= len[b] the output is 0 and (c, h) are carried.
    Backward direction = reverse_sequence over the first len[b] steps, run, reverse back.
    In-tree cross-reference of the cell equations: chiron/chiron/utils/lstm.py:37-46.
    """
    B, T, D = v.shape
    H = kernel.shape[1] // 4
    dt = v.dtype
    assert kernel.shape[0] == D + H, (kernel.shape, D, H)
    wx, wh = kernel[:D].astype(dt), kernel[D:].astype(dt)
    bias = bias.astype(dt)
    fb = dt.type(forget_bias)
    out = np.zeros((B, T, H), dtype=dt)
    h = np.zeros((B, H), dtype=dt)
    c = np.zeros((B, H), dtype=dt)
    lens = np.asarray(lens).astype(np.int64)
    maxlen = int(lens.max()) if B else 0
    rows = np.arange(B)
    for s in range(maxlen):
        active = s < lens
        t = np.where(reverse, lens - 1 - s, s)
        t = np.where(active, t, 0)
        x_t = v[rows, t]                                   # [B, D]
        z = np.concatenate([x_t, h], axis=1) @ np.concatenate([wx, wh], axis=0) + bias
        i, j, f, o = z[:, :H], z[:, H:2 * H], z[:, 2 * H:3 * H], z[:, 3 * H:]
        c_new = sigmoid(f + fb) * c + sigmoid(i) * np.tanh(j)
        h_new = sigmoid(o) * np.tanh(c_new)
        a = active[:, None]
        c = np.where(a, c_new, c)
        h = np.where(a, h_new, h)
        idx = np.nonzero(active)[0]
        out[idx, t[idx]] = h_new[idx]
    return out


def blstm(v, lens, w, forget_bias):
    """fw then bw concatenated on the feature axis (network.py:285 ``tf.concat(outputs, 2)``)."""
    fw = lstm_dir(v, lens, w["kernel_fw"], w["bias_fw"], forget_bias, False)
    bw = lstm_dir(v, lens, w["kernel_bw"], w["bias_bw"], forget_bias, True)
    return np.concatenate([fw, bw], axis=2)


# ------------------------------------------------------------------ a4: residual layer (network.py:312-349, 385-400)
def conv(v, W):
    """tf.nn.conv2d NHWC on [B,1,T,Cin], filter [1,k,Cin,Cout], stride 1, SAME, no bias
    (network.py:327-329): cross-correlation with (k-1)//2 zeros on the left."""
    B, T, Cin = v.shape
    k, cin, cout = W.shape[1], W.shape[2], W.shape[3]
    assert cin == Cin
    dt = v.dtype
    W = W.astype(dt)
    left = (k - 1) // 2
    out = np.zeros((B, T, cout), dtype=dt)
    flat = None
    for dk in range(k):
        sh = dk - left
        src = np.zeros_like(v)
        if sh == 0:
            src = v
        elif sh < 0:
            src[:, -sh:] = v[:, :T + sh]
        else:
            src[:, :T - sh] = v[:, sh:]
        flat = src.reshape(B * T, Cin) @ W[0, dk]
        out += flat.reshape(B, T, cout)
    return out


def bn(y, scale, offset, eps=1e-5):
    """simple_global_bn (network.py:339-349): tf.nn.moments over axes [0,1,2] of THIS batch
    (also at inference, padded positions included), then tf.nn.batch_normalization:
    inv = rsqrt(var + eps) * scale ; y * inv + (offset - mean * inv)."""
    dt = y.dtype
    flat = y.reshape(-1, y.shape[-1])
    mean = flat.mean(axis=0, dtype=dt)
    var = np.mean(np.square(flat - mean), axis=0, dtype=dt)
    inv = (dt.type(1) / np.sqrt(var + dt.type(eps))) * scale.astype(dt)
    return y * inv + (offset.astype(dt) - mean * inv)


def residual_layer(v, w, batchnorm):
    """network.py:385-400.  branch1: conv1x1 (+BN iff ``batchnorm``, no ReLU);
    branch2: conv1x1+BN+ReLU -> conv1x3+BN+ReLU -> conv1x1+BN ; relu(branch1 + branch2)."""
    a = conv(v, w["conv1"])
    if batchnorm:
        a = bn(a, w["conv1_bn_scale"], w["conv1_bn_offset"])
    b = np.maximum(bn(conv(v, w["conv2a"]), w["conv2a_bn_scale"], w["conv2a_bn_offset"]), 0)
    b = np.maximum(bn(conv(b, w["conv2b"]), w["conv2b_bn_scale"], w["conv2b_bn_offset"]), 0)
    b = bn(conv(b, w["conv2c"]), w["conv2c_bn_scale"], w["conv2c_bn_offset"])
    return np.maximum(a + b, 0)


# ------------------------------------------------------------------ a5: chiron_fnn (network.py:164-175)
def chiron_fnn(v, w):
    H = v.shape[2] // 2
    dt = v.dtype
    return v[:, :, :H] * w["weights"][0].astype(dt) + v[:, :, H:] * w["weights"][1].astype(dt) \
        + w["bias"].astype(dt)


# ------------------------------------------------------------------ a7 / a8: predictions
def argmax_first(logits):
    """tf.argmax: first index on ties, int64 (network.py:208-211)."""
    return np.argmax(logits, axis=-1).astype(np.int64)


def ctc_greedy(logits, lens):
    """tf.nn.ctc_greedy_decoder(merge_repeated=True) + sparse_tensor_to_dense(default -1)
    (network.py:528-533).  blank = C-1; prev is updated at every step, blanks included.
    Returns (dense int64 [B, max_decoded], neg_sum_logits [B])."""
    B, T, C = logits.shape
    best = np.argmax(logits, axis=2)
    seqs = []
    neg = np.zeros((B,), dtype=logits.dtype)
    for b in range(B):
        prev = -1
        s = []
        for t in range(int(lens[b])):
            k = int(best[b, t])
            neg[b] += -logits[b, t, k]
            if k != C - 1 and k != prev:
                s.append(k)
            prev = k
        seqs.append(s)
    width = max((len(s) for s in seqs), default=0)
    dense = np.full((B, width), -1, dtype=np.int64)
    for b, s in enumerate(seqs):
        dense[b, :len(s)] = s
    return dense, neg


# ------------------------------------------------------------------ a1/a2/a6: the graph
def forward_batch(x, lens, network, weights, graph, dtype=np.float32, return_trace=False):
    """ONE reference batch (= one BN group).  x: [B,T] (CtcLoss) or [B,T,len_x] (CrossEntropy).

    Returns (logits, labels): CrossEntropy logits [B*T, C], labels int64 [B*T];
    CtcLoss logits [B, T, C], labels int64 [B, max_decoded] padded with -1."""
    x = np.asarray(x, dtype=dtype)
    lens = np.asarray(lens, dtype=np.int32)
    if graph == "CtcLoss":
        v = x.reshape(x.shape[0], x.shape[1], 1)           # network.py:515-518
    elif graph == "CrossEntropy":
        v = x
    else:
        raise ValueError(graph)
    trace = []
    for layer, w in zip(network, weights["layers"]):
        t = layer["type"]
        if t == "residual_layer":
            v = residual_layer(v, w, bool(layer.get("batchnorm")))
        elif t == "blstm":
            v = blstm(v, lens, w, layer["bias"])
        elif t == "chiron_fnn":
            v = chiron_fnn(v, w)
        else:
            raise ValueError("layer type %r is not on the supported forward path" % t)
        if return_trace:
            trace.append(v)
    B, T, F = v.shape
    logits = v.reshape(B * T, F) @ weights["out_w"].astype(dtype) + weights["out_b"].astype(dtype)
    if graph == "CrossEntropy":
        labels = argmax_first(logits)
    else:
        logits = logits.reshape(B, T, -1)
        labels, _ = ctc_greedy(logits, lens)
    if return_trace:
        return logits, labels, trace
    return logits, labels


def forward(x, lens, network, weights, graph, batch_size, dtype=np.float32):
    """The reference's inner loop (run_nanotensor.py:343-347): consecutive ``batch_size`` windows
    per sess.run, last batch short.  Returns a list of per-batch (logits, labels)."""
    out = []
    n = len(x)
    for s in range(0, n, batch_size):
        out.append(forward_batch(x[s:s + batch_size], lens[s:s + batch_size], network, weights,
                                 graph, dtype))
    return out


# ------------------------------------------------------------------ a9: signal batcher
def normalise_read(raw):
    """chiron_input.read_signal(normalize=True) (chiron/chiron/chiron_input.py:220-230):
    z-score with the mean / population std of the whole read, in float64."""
    raw = np.asarray(raw, dtype=np.float64)
    return (raw - np.mean(raw)) / np.std(raw)


def windows_of_read(raw, seq_len, step=300, start_index=0, normalise=True):
    """FullSignalSequence.load_data_inference (dataset.py:435-445): windows every ``step``
    samples, zero padded to ``seq_len``; yields float32 x [n, seq_len], int32 len [n]."""
    s = normalise_read(raw) if normalise else np.asarray(raw, dtype=np.float64)
    s = s[start_index:]
    n = len(s)
    starts = list(range(0, n, step))
    x = np.zeros((len(starts), seq_len), dtype=np.float32)
    lens = np.zeros((len(starts),), dtype=np.int32)
    for i, st in enumerate(starts):
        seg = s[st:st + seq_len]
        x[i, :len(seg)] = seg                               # float64 -> float32 at the tf feed
        lens[i] = len(seg)
    return x, lens


# ------------------------------------------------------------------ a10: event batcher
def event_windows(data, seq_len):
    """NumpyEventData.load_data (dataset.py:490-517) for one ``.npy`` object array [n,2]:
    len(data)//seq_len full windows, tail dropped."""
    nb = len(data) // seq_len
    xs, ys = [], []
    for b in range(nb):
        chunk = data[b * seq_len:(b + 1) * seq_len]
        xs.append(np.vstack(chunk[:, 0]))
        ys.append(np.vstack(chunk[:, 1]))
    return np.asarray(xs), np.asarray(ys), np.full((nb,), seq_len, dtype=np.int32)


# ------------------------------------------------------------------ f1: post-processing
def index2base(read):
    """SignalLabel.index2base (nanotensor/trim_signal.py:116-122), blank=False."""
    base = ["A", "C", "G", "T", "E"]
    return "".join(base[int(k)] if int(k) != -1 else "" for k in read)


def simple_assembly(bpreads):
    """chiron/chiron/utils/easy_assembler.py:11-37 (difflib longest matching block placement)."""
    consensus = np.zeros([5, 1000])
    pos = 0
    length = 0
    census_len = 1000
    lut = {"A": 0, "C": 1, "G": 2, "T": 3, "E": 4}

    def add(start, seg):
        if start < 0:
            seg = seg[-start:]
            start = 0
        for i, ch in enumerate(seg):
            consensus[lut[ch.upper()]][start + i] += 1

    for indx, bpread in enumerate(bpreads):
        if indx == 0:
            add(0, bpread)
            continue
        d = difflib.SequenceMatcher(None, bpreads[indx - 1], bpread)
        mb = max(d.get_matching_blocks(), key=lambda m: m[2])
        disp = mb[0] - mb[1]
        if disp + pos + len(bpread) > census_len:
            consensus = np.pad(consensus, ((0, 0), (0, 1000)), mode="constant")
            census_len += 1000
        add(pos + disp, bpread)
        pos += disp
        length = max(length, pos + len(bpread))
    return consensus[:, :length]


def process_output(label_batches):
    """dataset.py:421-432: per-window label rows -> consensus read string."""
    reads = []
    for batch in label_batches:
        reads.extend(index2base(r) for r in batch)
    cons = simple_assembly(reads)
    return index2base(np.argmax(cons, axis=0)) if cons.shape[1] else ""
