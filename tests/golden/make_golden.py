#!/usr/bin/env python
"""Generates the fixtures under tests/golden/ (run in the build container, where /root/reference
exists; the GPU box only reads the committed outputs).

Inputs come from the REFERENCE's shipped data:
  test_files/test_models/08Aug-25-16h-35m/test_model-30.{index,data-00000-of-00001}  (weights)
  test_files/create_training_files/07Jul-20-11h-28m/nanonet_canonical{0..4}.npy       (event windows)
Outputs are the CPU oracle's (oracle/oracle_np.py) logits / labels on them, cross-checked here
against the independent torch oracle.  The reference ships NO golden outputs for this path
("parity unpinned", see oracle/oracle_np.py), so these are OUR goldens on ITS inputs.

Also writes small synthetic CRNN cases (weights are regenerated from seeds by
nanopore_rnn_b200.synthetic.random_weights, inputs and oracle outputs are stored).
"""
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import nanopore_rnn_b200  # noqa: E402,F401
from nanopore_rnn_b200 import checkpoint as ckpt  # noqa: E402
from nanopore_rnn_b200 import synthetic as syn  # noqa: E402
import oracle_np as A  # noqa: E402
import oracle_torch as B  # noqa: E402

REF = "/root/reference"
CKPT = os.path.join(REF, "test_files/test_models/08Aug-25-16h-35m/test_model-30")
NPY = os.path.join(REF, "test_files/create_training_files/07Jul-20-11h-28m/nanonet_canonical%d.npy")


def event_fixture():
    for ext in (".index", ".data-00000-of-00001"):
        shutil.copyfile(CKPT + ext, os.path.join(HERE, "test_model-30" + ext))
    for p in (os.path.join(HERE, "test_model-30.index"), os.path.join(HERE, "test_model-30.data-00000-of-00001")):
        os.chmod(p, 0o644)
    xs, ys = [], []
    for i in range(5):
        data = np.load(NPY % i, allow_pickle=True, encoding="latin1")
        x, y, _ = A.event_windows(data, 30)
        xs.append(x)
        ys.append(y.argmax(axis=2))
    x = np.concatenate(xs).astype(np.float32)          # [80, 30, 12]
    y = np.concatenate(ys).astype(np.int16)            # [80, 30] class indices of the one-hot labels
    lens = np.full((80,), 30, np.int32)
    net = syn.NETWORKS["event_fixture"]
    w = ckpt.weights_for_network(ckpt.load_checkpoint(CKPT), net)
    logits, labels = A.forward_batch(x, lens, net, w, "CrossEntropy")
    lt = B.forward_batch(x, lens, net, w, "CrossEntropy")
    l64, _ = A.forward_batch(x, lens, net, w, "CrossEntropy", dtype=np.float64)
    print("event fixture: np-vs-torch %.2e  np-vs-fp64 %.2e  acc %.4f" % (
        np.abs(logits - lt).max(), np.abs(logits - l64).max(), float((labels == y.reshape(-1)).mean())))
    assert np.abs(logits - lt).max() < 1e-5
    np.savez_compressed(os.path.join(HERE, "event_fixture.npz"), x=x, y=y, lens=lens,
                        logits_head=logits[:8 * 30].astype(np.float32),      # first 8 windows
                        logits_rowmax=logits.max(axis=1).astype(np.float32),
                        labels=labels.astype(np.int16),
                        accuracy=np.float64((labels == y.reshape(-1)).mean()))


def crnn_case(name, net_key, n_classes, batch, n_reads, wpr, seed):
    net = syn.NETWORKS[net_key]
    w = syn.random_weights(net, 1, n_classes, seed=seed + 2, readout_gain=1.0)
    x, lens, rid = syn.synthetic_reads(n_reads, wpr, 100, seed=seed)
    lens[3] = 0                                         # an empty window (seq_len 0)
    x[3] = 0
    goff = []
    from nanopore_rnn_b200.engine import group_offsets
    goff = group_offsets([wpr] * n_reads, batch)
    logits, best = [], []
    worst = 0.0
    for g in range(len(goff) - 1):
        a, b = goff[g], goff[g + 1]
        lg, _ = A.forward_batch(x[a:b], lens[a:b], net, w, "CtcLoss")
        lt = B.forward_batch(x[a:b], lens[a:b], net, w, "CtcLoss")
        worst = max(worst, float(np.abs(lg - lt).max()))
        logits.append(lg)
    logits = np.concatenate(logits)
    labels, _ = A.ctc_greedy(logits, lens)
    T = logits.shape[1]
    dense = np.full((len(x), T), -1, np.int16)
    dense[:, :labels.shape[1]] = labels
    hist = np.bincount(logits.argmax(2).reshape(-1), minlength=n_classes)
    print("%s: np-vs-torch %.2e, argmax histogram %s, decoded width %d" % (name, worst, hist, labels.shape[1]))
    assert worst < 5e-4
    np.savez_compressed(os.path.join(HERE, name + ".npz"), x=x, lens=lens, group_off=goff,
                        logits=logits.astype(np.float32), labels=dense, seed=np.int64(seed),
                        n_classes=np.int64(n_classes), batch=np.int64(batch))


if __name__ == "__main__":
    event_fixture()
    crnn_case("crnn_signal", "signal_CRNN", 6, 20, 2, 25, seed=10)       # groups 20,5,20,5
    crnn_case("crnn_run_nanotensor", "run_nanotensor", 4, 10, 2, 13, seed=0)   # groups 10,3,10,3
