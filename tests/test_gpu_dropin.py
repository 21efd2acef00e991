"""GPU (-m gpu): the reference-facing Python surface (CtcLoss / CrossEntropy / RunTensorflow)
driven the way the reference's commented-out tests drove it (nanotensor/tests/network_test.py,
run_nanotensor_test.py), on the B200 path."""
import json
import os

import numpy as np
import pytest

import oracle_np as A
from nanopore_rnn_b200 import dataset, network, run_nanotensor, synthetic as syn

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cross_entropy_graph_on_shipped_checkpoint(golden_dir):
    g = np.load(os.path.join(golden_dir, "event_fixture.npz"))
    net = syn.NETWORKS["event_fixture"]
    graph = network.CrossEntropy(network=net, x_shape=[None, 30, 12], seq_shape=[None], y_shape=[None, 30, 1025],
                                 len_x=12, len_y=1025, seq_len=30, mode=1,
                                 model_path=os.path.join(golden_dir, "test_model-30"))
    pred, out, bs = graph.run_batch(g["x"][:7], g["lens"][:7],
                                    fetches=[graph.prediction, graph.output_layer, graph.batch_size])
    assert bs == 7 and pred.dtype == np.int64 and pred.shape == (7 * 30,) and out.shape == (7 * 30, 1025)
    assert np.abs(out - g["logits_head"][:210]).max() <= 1e-3
    assert np.array_equal(pred, g["labels"][:210].astype(np.int64))
    acc = graph.run_batch(g["x"], g["lens"], fetches=graph.accuracy, y=g["y"])
    assert abs(acc - float(g["accuracy"])) < 1e-9


def test_ctc_graph_prediction_shape_conventions():
    net = syn.NETWORKS["signal_CRNN"]
    w = syn.random_weights(net, 1, 6, seed=12, readout_gain=8.0)
    x, lens, _ = syn.synthetic_reads(1, 20, 100, seed=11)
    graph = network.CtcLoss(network=net, x_shape=[None, 100], seq_shape=[None], len_x=1, len_y=5, seq_len=100,
                            mode=2, weights=w)
    assert graph.len_y == 6                                    # blank appended (network.py:495)
    pred, out = graph.run_batch(x, lens, fetches=[graph.prediction, graph.output_layer])
    ref_logits, ref_pred = A.forward_batch(x, lens, net, w, "CtcLoss")
    assert out.shape == (20, 100, 6) and pred.dtype == np.int64
    assert np.abs(out - ref_logits).max() <= 8e-3
    if np.array_equal(out.argmax(2), ref_logits.argmax(2)):
        assert pred.shape == ref_pred.shape and np.array_equal(pred, ref_pred)   # [B, max_decoded], pad -1
    # the decoder's second output (dropped by the reference, network.py:531) and Chiron's beam-search prediction
    import oracle_ctc as O
    neg, (dec, logp) = graph.run_batch(x, lens, fetches=[graph.neg_sum_logits, graph.beam_prediction])
    _, _, rneg = O.ctc_greedy(out, lens)
    assert neg.shape == (20, 1) and np.array_equal(neg[:, 0], rneg)
    rp, rl, rlp = O.ctc_beam_search(out[:4], lens[:4], 100, 1, False)
    assert np.allclose(logp[:4], rlp, atol=1e-3)
    for b in range(4):
        assert [int(v) for v in dec[0][b] if v >= 0] == [int(v) for v in rp[b, 0, :rl[b, 0]]]


def test_run_tensorflow_inference_and_test_routes(golden_dir, tmp_path):
    # inference route on .signal files with a CRNN checkpoint written in the TF bundle format
    from nanopore_rnn_b200 import checkpoint as ckpt
    rng = np.random.default_rng(0)
    sig_dir = tmp_path / "reads"
    sig_dir.mkdir()
    for i, n in enumerate([2300, 640]):
        (sig_dir / ("r%d.signal" % i)).write_text(" ".join(str(int(v)) for v in rng.integers(300, 700, n)))
    cfg = json.load(open(os.path.join(golden_dir, "configs", "signal_CRNN.config.json")))
    net = cfg["network"]
    w = syn.random_weights(net, 1, 6, seed=21, readout_gain=8.0)
    prefix = str(tmp_path / "small_test-1")
    ckpt.save_checkpoint(prefix, ckpt.variables_from_weights(w, net, model_name=cfg["model_name"]))
    cfg.update(train=False, inference=True, test=False, use_checkpoint=False, trained_model_path=prefix)
    cfg["CreateDataset"]["inference_dir"] = str(sig_dir)
    cfg["CreateDataset"]["dataset_args"].update(inference_output_dir=str(tmp_path), step=100)
    args = run_nanotensor.CommandLine.check_args(cfg)
    runner = run_nanotensor.RunTensorflow(args)
    outs = runner.run_model()
    outs_pb = run_nanotensor.RunTensorflow(args).run_model(per_batch=True)
    assert [open(o).read() for o in outs] == [open(o).read() for o in outs_pb]
    # oracle route for read 0
    ds = dataset.FullSignalSequence([str(sig_dir / "r0.signal")], mode=2, batch_size=20, seq_len=100, step=100,
                                    verbose=False)
    x, lens, goff, _ = ds.inference_job()
    batches = [A.forward_batch(x[a:b], lens[a:b], net, w, "CtcLoss")[1] for a, b in zip(goff[:-1], goff[1:])]
    fasta = open(outs[0]).read().split("\n")[1]
    # exact comparison against the oracle route: the FASTA must equal the oracle's consensus whenever every argmax
    # agrees; positions that flip must be near ties in the oracle (readout gain 8 scales the 1e-3 logits budget), and
    # then the FASTA must still equal the oracle's assembly of the GPU's own decoded labels
    from nanopore_rnn_b200 import engine
    labels, llen, logits = runner.inference_model.model.forward_host(x, lens, goff, want_logits=True)
    ref_logits = np.concatenate([A.forward_batch(x[a:b], lens[a:b], net, w, "CtcLoss")[0]
                                 for a, b in zip(goff[:-1], goff[1:])]).reshape(-1, 6)
    assert np.abs(logits - ref_logits).max() <= 8e-3
    flips = np.argwhere(logits.argmax(1) != ref_logits.argmax(1)).reshape(-1)
    for r in flips:
        top = np.sort(ref_logits[r])[-2:]
        assert top[1] - top[0] <= 2 * 8e-3, (int(r), float(top[1] - top[0]))
    assert len(flips) <= 1e-3 * len(ref_logits)
    gpu_batches = [engine.ctc_dense(labels[a:b], llen[a:b]) for a, b in zip(goff[:-1], goff[1:])]
    assert fasta == A.process_output(gpu_batches)
    if len(flips) == 0:
        assert fasta == A.process_output(batches)
    assert len(fasta) >= 10 and len(set(fasta)) >= 2         # and the case is not degenerate (19 bases, 3 letters)

    # test route: CrossEntropy accuracy on the shipped event checkpoint
    arr_dir = tmp_path / "npy"
    arr_dir.mkdir()
    g = np.load(os.path.join(golden_dir, "event_fixture.npz"))
    obj = np.empty((80 * 30, 2), dtype=object)
    flat_x, flat_y = g["x"].reshape(-1, 12), g["y"].reshape(-1)
    for i in range(len(obj)):
        obj[i, 0] = flat_x[i].astype(np.float64)
        one = np.zeros(1025)
        one[flat_y[i]] = 1
        obj[i, 1] = one
    np.save(str(arr_dir / "a.npy"), obj, allow_pickle=True)
    tcfg = {"CreateDataset": {"dataset": "NumpyEventData", "test_dir": str(arr_dir),
                              "dataset_args": {"batch_size": 1, "seq_len": 30}},
            "BuildGraph": {"graph": "CrossEntropy", "graph_args": {}},
            "network": syn.NETWORKS["event_fixture"], "train": False, "inference": False, "test": True,
            "use_checkpoint": True, "trained_model": golden_dir, "model_name": "my_test_model"}
    acc = run_nanotensor.RunTensorflow(run_nanotensor.CommandLine.check_args(tcfg)).test_model()
    assert abs(acc - 100 * float(g["accuracy"])) < 1e-6
