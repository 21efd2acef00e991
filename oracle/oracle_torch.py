"""SECOND CPU ORACLE (torch) -- TEST INFRASTRUCTURE ONLY; see oracle/oracle_np.py's header
("parity unpinned").  An independent build of the same forward out of library ops, so that a
slip in the hand restatement shows up as a disagreement between the two:

* LSTM  -> torch.nn.LSTM(bidirectional) with TF's gate columns i|j|f|o permuted to torch's
           i|f|g|o rows, forget_bias folded into b_ih's f block, b_hh = 0, and
           pack_padded_sequence for dynamic_rnn's masking / reverse_sequence semantics
           (network.py:266-288);
* conv  -> F.conv1d with padding=(k-1)//2 (network.py:312-329);
* BN    -> batch moments over (B,T) per channel, biased variance (network.py:339-349).
"""
import numpy as np
import torch
import torch.nn.functional as F
from torch.nn.utils.rnn import pack_padded_sequence, pad_packed_sequence


def _perm_gates(a, H):
    # TF blocks (i, j, f, o) -> torch blocks (i, f, g=j, o) along the leading axis
    i, j, f, o = a[:H], a[H:2 * H], a[2 * H:3 * H], a[3 * H:]
    return torch.cat([i, f, j, o], 0)


def blstm(v, lens, w, forget_bias):
    B, T, D = v.shape
    H = w["kernel_fw"].shape[1] // 4
    m = torch.nn.LSTM(D, H, batch_first=True, bidirectional=True).to(v.dtype)
    with torch.no_grad():
        for d, sfx in (("fw", ""), ("bw", "_reverse")):
            k = torch.as_tensor(w["kernel_" + d], dtype=v.dtype)
            b = torch.as_tensor(w["bias_" + d], dtype=v.dtype).clone()
            b[2 * H:3 * H] += forget_bias
            getattr(m, "weight_ih_l0" + sfx).copy_(_perm_gates(k[:D].t().contiguous(), H))
            getattr(m, "weight_hh_l0" + sfx).copy_(_perm_gates(k[D:].t().contiguous(), H))
            getattr(m, "bias_ih_l0" + sfx).copy_(_perm_gates(b, H))
            getattr(m, "bias_hh_l0" + sfx).zero_()
    lens_t = torch.as_tensor(np.asarray(lens), dtype=torch.int64)
    out = torch.zeros(B, T, 2 * H, dtype=v.dtype)
    nz = torch.nonzero(lens_t > 0).flatten()
    if len(nz):
        packed = pack_padded_sequence(v[nz], lens_t[nz], batch_first=True, enforce_sorted=False)
        with torch.no_grad():
            y, _ = m(packed)
        y, _ = pad_packed_sequence(y, batch_first=True, total_length=T)
        out[nz] = y
    return out


def conv(v, W):
    Wt = torch.as_tensor(W, dtype=v.dtype)[0]          # [k, Cin, Cout]
    k = Wt.shape[0]
    return F.conv1d(v.transpose(1, 2), Wt.permute(2, 1, 0).contiguous(), padding=(k - 1) // 2).transpose(1, 2)


def bn(y, scale, offset, eps=1e-5):
    mean = y.mean(dim=(0, 1))
    var = y.var(dim=(0, 1), unbiased=False)
    return (y - mean) / torch.sqrt(var + eps) * torch.as_tensor(scale, dtype=y.dtype) \
        + torch.as_tensor(offset, dtype=y.dtype)


def residual_layer(v, w, batchnorm):
    a = conv(v, w["conv1"])
    if batchnorm:
        a = bn(a, w["conv1_bn_scale"], w["conv1_bn_offset"])
    b = F.relu(bn(conv(v, w["conv2a"]), w["conv2a_bn_scale"], w["conv2a_bn_offset"]))
    b = F.relu(bn(conv(b, w["conv2b"]), w["conv2b_bn_scale"], w["conv2b_bn_offset"]))
    b = bn(conv(b, w["conv2c"]), w["conv2c_bn_scale"], w["conv2c_bn_offset"])
    return F.relu(a + b)


def forward_batch(x, lens, network, weights, graph, dtype=torch.float32):
    """Logits only (labels are derived by the numpy oracle's decoders)."""
    v = torch.as_tensor(np.asarray(x), dtype=dtype)
    if graph == "CtcLoss":
        v = v.reshape(v.shape[0], v.shape[1], 1)
    for layer, w in zip(network, weights["layers"]):
        t = layer["type"]
        if t == "residual_layer":
            v = residual_layer(v, w, bool(layer.get("batchnorm")))
        elif t == "blstm":
            v = blstm(v, lens, w, layer["bias"])
        elif t == "chiron_fnn":
            H = v.shape[2] // 2
            fw = torch.as_tensor(w["weights"], dtype=dtype)
            v = v[:, :, :H] * fw[0] + v[:, :, H:] * fw[1] + torch.as_tensor(w["bias"], dtype=dtype)
        else:
            raise ValueError(t)
    B, T, Fdim = v.shape
    logits = v.reshape(B * T, Fdim) @ torch.as_tensor(weights["out_w"], dtype=dtype) \
        + torch.as_tensor(weights["out_b"], dtype=dtype)
    if graph == "CtcLoss":
        logits = logits.reshape(B, T, -1)
    return logits.numpy()


# ------------------------------------------------------------------ many reference batches at once
def _bn_groups(y, scale, offset, group, eps=1e-5):
    """bn() applied independently to consecutive groups of ``group`` windows (one reference batch
    each, network.py:339-349), vectorised over the groups."""
    n, T, C = y.shape
    g = y.reshape(n // group, group * T, C)
    mean = g.mean(dim=1, keepdim=True)
    var = g.var(dim=1, unbiased=False, keepdim=True)
    out = (g - mean) / torch.sqrt(var + eps) * torch.as_tensor(scale, dtype=y.dtype) \
        + torch.as_tensor(offset, dtype=y.dtype)
    return out.reshape(n, T, C)


def forward_job(x, lens, network, weights, graph, group, dtype=torch.float32):
    """Logits of MANY equal-sized reference batches in one pass: the CPU's best case (large GEMMs,
    all threads) with the reference's per-batch BN statistics kept.  ``len(x)`` must be a multiple of
    ``group`` (= the config's batch_size).  Used as the timed CPU baseline (bench.py) and checked
    against forward_batch / oracle_np in tests/test_oracle.py."""
    v = torch.as_tensor(np.asarray(x), dtype=dtype)
    assert v.shape[0] % group == 0
    if graph == "CtcLoss":
        v = v.reshape(v.shape[0], v.shape[1], 1)

    def gbn(y, w, name):
        return _bn_groups(y, w[name + "_bn_scale"], w[name + "_bn_offset"], group)

    with torch.no_grad():
        for layer, w in zip(network, weights["layers"]):
            t = layer["type"]
            if t == "residual_layer":
                a = conv(v, w["conv1"])
                if layer.get("batchnorm"):
                    a = gbn(a, w, "conv1")
                b = F.relu(gbn(conv(v, w["conv2a"]), w, "conv2a"))
                b = F.relu(gbn(conv(b, w["conv2b"]), w, "conv2b"))
                b = gbn(conv(b, w["conv2c"]), w, "conv2c")
                v = F.relu(a + b)
            elif t == "blstm":
                v = blstm(v, lens, w, layer["bias"])
            elif t == "chiron_fnn":
                H = v.shape[2] // 2
                fw = torch.as_tensor(w["weights"], dtype=dtype)
                v = v[:, :, :H] * fw[0] + v[:, :, H:] * fw[1] + torch.as_tensor(w["bias"], dtype=dtype)
            else:
                raise ValueError(t)
        B, T, Fdim = v.shape
        logits = v.reshape(B * T, Fdim) @ torch.as_tensor(weights["out_w"], dtype=dtype) \
            + torch.as_tensor(weights["out_b"], dtype=dtype)
        if graph == "CtcLoss":
            logits = logits.reshape(B, T, -1)
    return logits.numpy()
