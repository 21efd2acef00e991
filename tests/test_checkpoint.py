"""CPU: TF-bundle reader against the checkpoint the reference ships, and name normalisation."""
import os

import numpy as np
import pytest

from nanopore_rnn_b200 import checkpoint as ckpt
from nanopore_rnn_b200 import synthetic as syn


def test_shipped_checkpoint_variable_table(golden_dir):
    prefix = os.path.join(golden_dir, "test_model-30")
    table = dict(ckpt.list_variables(prefix))
    assert table["blstm_layer1/bidirectional_rnn/fw/lstm_cell/weights"] == (17, 20)
    assert table["blstm_layer1/bidirectional_rnn/bw/lstm_cell/biases"] == (20,)
    assert table["weights"] == (10, 1025) and table["bias"] == (1025,)
    assert len(table) == 36
    v = ckpt.load_checkpoint(prefix, skip_optimizer=False)
    assert v["global_step"] == np.float32(30.0)
    assert v["global_step"].dtype == np.float32 and v["global_step_1"].dtype == np.int32
    assert np.isfinite(v["weights"]).all() and v["weights"].std() > 0


def test_weights_for_network_old_tf_names(golden_dir):
    v = ckpt.load_checkpoint(os.path.join(golden_dir, "test_model-30"))
    assert not any("Adam" in k for k in v)
    w = ckpt.weights_for_network(v, syn.NETWORKS["event_fixture"])
    assert w["layers"][0]["kernel_fw"].shape == (17, 20) and w["out_w"].shape == (10, 1025)
    # TF >= 1.2 names and a model_name scope resolve to the same tensors
    v2 = {}
    for k, a in v.items():
        k2 = k.replace("lstm_cell/weights", "lstm_cell/kernel").replace("lstm_cell/biases", "lstm_cell/bias")
        v2["my_test_model/" + k2] = a
    w2 = ckpt.weights_for_network(v2, syn.NETWORKS["event_fixture"], model_name="my_test_model")
    np.testing.assert_array_equal(w2["layers"][0]["kernel_bw"], w["layers"][0]["kernel_bw"])
    w3 = ckpt.weights_for_network(v2, syn.NETWORKS["event_fixture"])          # prefix guessed
    np.testing.assert_array_equal(w3["out_b"], w["out_b"])


def test_missing_variable_and_bad_files(golden_dir, tmp_path):
    v = ckpt.load_checkpoint(os.path.join(golden_dir, "test_model-30"))
    with pytest.raises(KeyError):
        ckpt.weights_for_network(v, syn.NETWORKS["signal_CRNN"])
    bad = tmp_path / "x.index"
    bad.write_bytes(b"\x00" * 100)
    with pytest.raises(ckpt.CheckpointError):
        ckpt.read_index(str(bad))
    bad.write_bytes(b"\x00" * 10)
    with pytest.raises(ckpt.CheckpointError):
        ckpt.read_index(str(bad))


def test_latest_checkpoint(golden_dir, tmp_path):
    assert ckpt.latest_checkpoint(golden_dir).endswith("test_model-30")
    (tmp_path / "checkpoint").write_text('model_checkpoint_path: "m-7"\n')
    (tmp_path / "m-7.index").write_bytes(b"")
    (tmp_path / "m-3.index").write_bytes(b"")
    assert ckpt.latest_checkpoint(str(tmp_path)).endswith("m-7")


def test_writer_round_trip_and_crc(tmp_path):
    net = syn.NETWORKS["signal_CRNN"]
    w = syn.random_weights(net, 1, 6, seed=3)
    prefix = str(tmp_path / "small_test-5")
    ckpt.save_checkpoint(prefix, ckpt.variables_from_weights(w, net, model_name="small_test"))
    assert ckpt.latest_checkpoint(str(tmp_path)) == prefix
    table = dict(ckpt.list_variables(prefix))
    assert table["small_test/residual_layer_1/branch2/conv2b/weights"] == (1, 3, 256, 256)
    assert table["small_test/blstm_layer1/bidirectional_rnn/fw/lstm_cell/kernel"] == (356, 400)
    w2 = ckpt.weights_for_network(ckpt.load_checkpoint(prefix), net, model_name="small_test")
    for a, b in zip(w["layers"], w2["layers"]):
        assert sorted(a) == sorted(b)
        for k in a:
            np.testing.assert_array_equal(a[k], b[k])
    np.testing.assert_array_equal(w["out_w"], w2["out_w"])
    assert ckpt._crc32c(b"123456789") == 0xE3069283            # CRC-32C check value
    # the shipped file's block trailer verifies with the same CRC (format cross-check)
    import struct
    buf = open(os.path.join(os.path.dirname(__file__), "golden", "test_model-30.index"), "rb").read()
    footer = buf[-48:]
    pos = 0
    _, pos = ckpt._varint(footer, pos)
    _, pos = ckpt._varint(footer, pos)
    off, pos = ckpt._varint(footer, pos)
    size, pos = ckpt._varint(footer, pos)
    stored = struct.unpack_from("<I", buf, off + size + 1)[0]
    assert stored == ckpt._mask_crc(ckpt._crc32c(buf[off:off + size + 1]))
