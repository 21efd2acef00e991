"""CPU: the oracle against the committed goldens, the two oracles against each other, and the
TF semantics the restatement has to get right (masking, reverse, ties, BN, CTC)."""
import os

import numpy as np
import pytest

import oracle_np as A
import oracle_torch as B
from nanopore_rnn_b200 import checkpoint as ckpt
from nanopore_rnn_b200 import synthetic as syn
from nanopore_rnn_b200.engine import group_offsets


def _event(golden_dir):
    g = np.load(os.path.join(golden_dir, "event_fixture.npz"))
    net = syn.NETWORKS["event_fixture"]
    w = ckpt.weights_for_network(ckpt.load_checkpoint(os.path.join(golden_dir, "test_model-30")), net)
    return g, net, w


def test_event_fixture_matches_golden(golden_dir):
    g, net, w = _event(golden_dir)
    logits, labels = A.forward_batch(g["x"], g["lens"], net, w, "CrossEntropy")
    assert logits.shape == (80 * 30, 1025)
    np.testing.assert_allclose(logits[:240], g["logits_head"], atol=1e-6)
    np.testing.assert_allclose(logits.max(1), g["logits_rowmax"], atol=1e-6)
    assert np.array_equal(labels, g["labels"].astype(np.int64))
    assert abs(float((labels == g["y"].reshape(-1)).mean()) - float(g["accuracy"])) < 1e-12


def test_event_fixture_batching_invariant(golden_dir):
    # a pure-BLSTM graph has no cross-window coupling: B=1 batches == one batch of 80
    g, net, w = _event(golden_dir)
    whole, _ = A.forward_batch(g["x"][:6], g["lens"][:6], net, w, "CrossEntropy")
    parts = A.forward(g["x"][:6], g["lens"][:6], net, w, "CrossEntropy", batch_size=1)
    np.testing.assert_allclose(np.concatenate([p[0] for p in parts]), whole, atol=1e-6)


@pytest.mark.parametrize("name,key", [("crnn_signal", "signal_CRNN"), ("crnn_run_nanotensor", "run_nanotensor")])
def test_crnn_matches_golden_and_torch(golden_dir, name, key):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    net = syn.NETWORKS[key]
    C = int(g["n_classes"])
    w = syn.random_weights(net, 1, C, seed=int(g["seed"]) + 2)
    goff = g["group_off"]
    a, b = int(goff[1]), int(goff[2])          # the short group: cheapest
    lg, _ = A.forward_batch(g["x"][a:b], g["lens"][a:b], net, w, "CtcLoss")
    np.testing.assert_allclose(lg, g["logits"][a:b], atol=2e-5)
    lt = B.forward_batch(g["x"][a:b], g["lens"][a:b], net, w, "CtcLoss")
    np.testing.assert_allclose(lt, lg, atol=5e-4)
    labels, _ = A.ctc_greedy(g["logits"], g["lens"])
    dense = np.full((len(g["x"]), 100), -1)
    dense[:, :labels.shape[1]] = labels
    assert np.array_equal(dense, g["labels"])


def test_two_oracles_agree_blstm_ragged():
    rng = np.random.default_rng(5)
    net = [{"type": "blstm", "name": "l1", "size": 7, "bias": 5.0, "concat": True},
           {"type": "blstm", "name": "l2", "size": 6, "bias": 1.0, "concat": True}]
    w = syn.random_weights(net, 3, 9, seed=1, graph="CrossEntropy")
    x = rng.standard_normal((6, 11, 3)).astype(np.float32)
    lens = np.array([11, 0, 1, 5, 10, 11], np.int32)
    la, _ = A.forward_batch(x, lens, net, w, "CrossEntropy")
    lb = B.forward_batch(x, lens, net, w, "CrossEntropy")
    np.testing.assert_allclose(la, lb, atol=1e-5)


def test_lstm_masking_and_reverse_semantics():
    # dynamic_rnn: outputs at t >= len are exactly 0; the bw direction of a length-L row equals
    # the fw direction of the time-reversed first L steps, reversed back (reverse_sequence).
    rng = np.random.default_rng(0)
    D, H, T = 4, 3, 9
    k = rng.standard_normal((D + H, 4 * H)).astype(np.float32) * 0.5
    b = rng.standard_normal(4 * H).astype(np.float32) * 0.1
    x = rng.standard_normal((3, T, D)).astype(np.float32)
    lens = np.array([9, 4, 0])
    fw = A.lstm_dir(x, lens, k, b, 5.0, False)
    bw = A.lstm_dir(x, lens, k, b, 5.0, True)
    assert np.all(fw[1, 4:] == 0) and np.all(bw[1, 4:] == 0) and np.all(fw[2] == 0) and np.all(bw[2] == 0)
    xr = x.copy()
    xr[1, :4] = x[1, :4][::-1]
    xr[0] = x[0][::-1]
    fr = A.lstm_dir(xr, lens, k, b, 5.0, False)
    np.testing.assert_allclose(bw[1, :4], fr[1, :4][::-1], atol=1e-6)
    np.testing.assert_allclose(bw[0], fr[0][::-1], atol=1e-6)
    # garbage beyond len must not leak
    x2 = x.copy()
    x2[1, 4:] = 1e6
    np.testing.assert_array_equal(A.lstm_dir(x2, lens, k, b, 5.0, True)[1], bw[1])


def test_forget_bias_is_runtime_not_stored():
    rng = np.random.default_rng(1)
    k = rng.standard_normal((5, 8)).astype(np.float32)
    b = np.zeros(8, np.float32)
    x = rng.standard_normal((1, 4, 3)).astype(np.float32)
    a = A.lstm_dir(x, [4], k, b, 5.0, False)
    b2 = b.copy()
    b2[4:6] += 5.0                                   # f block = columns [2H, 3H)
    np.testing.assert_allclose(A.lstm_dir(x, [4], k, b2, 0.0, False), a, atol=1e-6)


def test_conv_same_padding_and_bn_batch_statistics():
    rng = np.random.default_rng(2)
    v = rng.standard_normal((2, 5, 3)).astype(np.float32)
    W = rng.standard_normal((1, 3, 3, 4)).astype(np.float32)
    y = A.conv(v, W)
    ref = np.zeros((2, 5, 4), np.float32)
    for t in range(5):
        for dk in range(3):
            s = t + dk - 1
            if 0 <= s < 5:
                ref[:, t] += v[:, s] @ W[0, dk]
    np.testing.assert_allclose(y, ref, atol=1e-5)
    out = A.bn(y, np.ones(4, np.float32), np.zeros(4, np.float32))
    flat = out.reshape(-1, 4)
    np.testing.assert_allclose(flat.mean(0), 0, atol=1e-5)
    np.testing.assert_allclose(flat.var(0), 1, atol=1e-3)
    # BN couples the windows of a batch: changing one window changes the others' outputs
    y2 = y.copy()
    y2[1] *= 3
    assert np.abs(A.bn(y2, np.ones(4, np.float32), np.zeros(4, np.float32))[0] - out[0]).max() > 1e-3


def test_argmax_first_and_ctc_greedy_rules():
    lg = np.array([[[1, 1, 0], [0, 2, 2], [0, 0, 0]]], np.float32)      # ties -> first index
    assert A.argmax_first(lg[0]).tolist() == [0, 1, 0]
    # blank = C-1 = 2; merge repeats; prev is updated on blanks too so "0 blank 0" -> 0 0
    seq = [0, 0, 2, 0, 1, 1, 2, 2, 1]
    lg = np.full((1, len(seq), 3), -1.0, np.float32)
    for t, k in enumerate(seq):
        lg[0, t, k] = 1.0
    dense, neg = A.ctc_greedy(lg, [len(seq)])
    assert dense.tolist() == [[0, 0, 1, 1]]
    assert neg[0] == -len(seq)
    dense, _ = A.ctc_greedy(lg, [2])
    assert dense.tolist() == [[0]]
    dense, _ = A.ctc_greedy(np.concatenate([lg, lg]), [0, 3])
    assert dense.tolist() == [[-1], [0]]
    assert A.ctc_greedy(lg, [0])[0].shape == (1, 0)


def test_window_batcher_semantics():
    raw = np.arange(1, 24, dtype=np.int64) * 3 % 17
    x, lens = A.windows_of_read(raw, seq_len=5, step=7, start_index=2)
    s = (raw - raw.mean()) / raw.std()
    s = s[2:]
    assert len(x) == len(range(0, len(s), 7))
    assert lens.tolist() == [5, 5, 5]
    np.testing.assert_allclose(x[1], s[7:12].astype(np.float32))
    x, lens = A.windows_of_read(raw, seq_len=10, step=10, start_index=0)
    assert lens.tolist() == [10, 10, 3] and np.all(x[2, 3:] == 0)


def test_group_offsets_follow_reference_batching():
    assert group_offsets([23, 10, 5], 10).tolist() == [0, 10, 20, 23, 33, 38]
    assert group_offsets([], 10).tolist() == [0]


def test_process_output_assembly():
    reads = ["AAGGCCTAGCT", "AGGCCTAGCAA", "GGCCTAGCTC", "AAAGGCCTAGT"]   # easy_assembler.py:40's own example
    cons = A.simple_assembly(reads)
    assert cons.shape[0] == 5 and cons.shape[1] > 0
    from nanopore_rnn_b200.dataset import simple_assembly, index2base
    np.testing.assert_array_equal(simple_assembly(reads), cons)
    assert index2base([0, 1, -1, 3, 4]) == "ACTE" == A.index2base([0, 1, -1, 3, 4])
    assert A.simple_assembly(["ACGT"]).shape == (5, 0)      # reference quirk: a single read has length 0
    np.testing.assert_array_equal(simple_assembly(["ACGT"]), A.simple_assembly(["ACGT"]))
    np.testing.assert_array_equal(simple_assembly(["ACGTAC", "TTACGTAC", "GTACC"]),
                                  A.simple_assembly(["ACGTAC", "TTACGTAC", "GTACC"]))


# ---------------------------------------------------------------- the C++/OpenMP restatement (oracle/oracle_cpu.cpp)
def _cpu():
    import subprocess
    import oracle_cpu as Cc
    if not os.path.exists(Cc.LIB_PATH):
        subprocess.check_call(["bash", os.path.join(os.path.dirname(Cc.LIB_PATH), "build.sh")])
    return Cc


def test_cpp_oracle_event_fixture_matches_golden(golden_dir):
    Cc = _cpu()
    g, net, w = _event(golden_dir)
    logits, labels, _ = Cc.forward(g["x"], g["lens"], net, w, "CrossEntropy", batch_size=16)
    np.testing.assert_allclose(logits[:240], g["logits_head"], atol=2e-5)
    np.testing.assert_allclose(logits.max(1), g["logits_rowmax"], atol=2e-5)
    assert float((labels == g["labels"]).mean()) >= 0.999


@pytest.mark.parametrize("name,key", [("crnn_signal", "signal_CRNN"), ("crnn_run_nanotensor", "run_nanotensor")])
def test_cpp_oracle_crnn_matches_golden(golden_dir, name, key):
    Cc = _cpu()
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    net = syn.NETWORKS[key]
    C = int(g["n_classes"])
    w = syn.random_weights(net, 1, C, seed=int(g["seed"]) + 2)
    logits, labels, llen = Cc.forward(g["x"], g["lens"], net, w, "CtcLoss", group_off=g["group_off"])
    np.testing.assert_allclose(logits, g["logits"].reshape(-1, C), atol=1e-4)
    # integer outputs from the GOLDEN logits are bit-exact; from its own logits they may differ only at near ties
    agree = float((labels == g["labels"]).mean())
    assert agree >= 0.999, agree


def test_cpp_oracle_ragged_blstm_and_ctc_rules():
    Cc = _cpu()
    rng = np.random.default_rng(9)
    net = [{"type": "blstm", "name": "l1", "size": 8, "bias": 5.0, "concat": True},
           {"type": "blstm", "name": "l2", "size": 8, "bias": 1.0, "concat": True}]
    w = syn.random_weights(net, 3, 5, seed=4, graph="CtcLoss", readout_gain=4.0)
    x = rng.standard_normal((9, 11, 3)).astype(np.float32)
    lens = np.array([11, 0, 1, 5, 11, 7, 2, 30, 10], dtype=np.int32)     # 0, 1, > T
    lg, labels, llen = Cc.forward(x, lens, net, w, "CtcLoss", batch_size=4)
    # the logits do not depend on the graph kind (oracle_np's CtcLoss graph insists on len_x = 1, network.py:515-518)
    ref, _ = A.forward_batch(x, np.minimum(lens, 11), net, w, "CrossEntropy")
    np.testing.assert_allclose(lg, ref.reshape(-1, 5), atol=2e-5)
    rl, _ = A.ctc_greedy(lg.reshape(9, 11, 5), np.minimum(lens, 11))
    dense = np.full((9, 11), -1)
    dense[:, :rl.shape[1]] = rl
    assert np.array_equal(labels, dense)
    assert np.array_equal(llen, (dense >= 0).sum(1))


def test_tensorflow_probe_reports_what_it_observes(golden_dir):
    """BASELINE.md section 2 / SURVEY.md section 7: the TF probe runs where the tests run.  Absent (expected): the
    outcome names the import error.  Present: the committed goldens must be reproduced by TF's own ops -- that pins
    every restatement under oracle/."""
    import tf_probe
    p = tf_probe.probe()
    assert set(p) == {"importable", "version", "error"}
    if not p["importable"]:
        assert p["error"] and "tensorflow" in tf_probe.describe().lower()
        return
    errs = tf_probe.pin_against_goldens(golden_dir)
    assert all(e <= 1e-4 for e in errs.values()), errs
