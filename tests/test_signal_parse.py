"""`.signal` text parse (chiron_input.py::read_signal's ``int(x) for x in line.split()``).

Parity PINNED: tests/golden/signal_text.npz = file bytes + what the reference's own read_signal returned
(tests/golden/make_golden_signal.py).  Integer work: oracle and CUDA path must reproduce the samples exactly.
"""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_signal as OS
from nanopore_rnn_b200 import _lib, dataset

from conftest import GOLDEN


@pytest.fixture(scope="module")
def golden():
    g = np.load(os.path.join(GOLDEN, "signal_text.npz"))
    return {k: g[k] for k in g.files}


def test_oracle_matches_reference_golden_exactly(golden):
    samples, off = OS.parse_signals(golden["text"].tobytes(), golden["file_off"])
    assert np.array_equal(off, golden["read_off"]) and np.array_equal(samples, golden["samples"])
    for bad in (b"12 1_000", b"12 3.5", b"40000", b"- 5", b"1234567"):
        with pytest.raises(ValueError):
            OS.parse_signal(bad)


@pytest.fixture(scope="module")
def lib():
    lib = _lib.load()
    assert lib.ntb_device_ok() == 1, lib.ntb_last_error()
    return lib


def _parse(lib, text, file_off, capacity=None):
    text = np.frombuffer(bytes(text), dtype=np.uint8)
    file_off = np.ascontiguousarray(file_off, dtype=np.int64)
    cap = len(text) // 2 + len(file_off) + 1 if capacity is None else capacity
    samples = np.full(max(cap, 1), -1, dtype=np.int16)
    off = np.zeros(len(file_off), dtype=np.int64)
    rc = lib.ntb_parse_signal_host(text.ctypes.data, file_off.ctypes.data, len(file_off) - 1, len(text), cap,
                                   samples.ctypes.data, off.ctypes.data, None)
    return rc, samples[:off[-1]], off


@pytest.mark.gpu
def test_gpu_matches_reference_golden_exactly(lib, golden):
    rc, samples, off = _parse(lib, golden["text"].tobytes(), golden["file_off"])
    assert rc == 0, lib.ntb_last_error()
    assert np.array_equal(off, golden["read_off"]) and np.array_equal(samples, golden["samples"])


@pytest.mark.gpu
def test_gpu_large_random_files_and_file_mirror(lib, tmp_path):
    """20 files of random length and formatting against the oracle; then through dataset.read_signals_gpu against
    dataset.read_signal_raw (the host parse) and into the z-score of read_signal."""
    rng = np.random.default_rng(8)
    seps = [" ", "\n", "\t", "  ", " \r\n"]
    blobs = []
    for f in range(20):
        vals = rng.integers(-32768, 32768, int(rng.integers(0, 60000)))
        blobs.append("".join(str(v) + seps[int(s)] for v, s in zip(vals, rng.integers(0, len(seps), len(vals)))).encode())
    off = np.concatenate([[0], np.cumsum([len(b) for b in blobs])])
    rc, samples, roff = _parse(lib, b"".join(blobs), off)
    want, woff = OS.parse_signals(b"".join(blobs), off)
    assert rc == 0 and np.array_equal(roff, woff) and np.array_equal(samples, want)
    paths = []
    for f, b in enumerate(blobs[:4]):
        p = tmp_path / ("r%d.signal" % f)
        p.write_bytes(b if len(b) else b"1 2 3")
        paths.append(str(p))
    got, goff = dataset.read_signals_gpu(paths)
    for k, p in enumerate(paths):
        assert np.array_equal(got[goff[k]:goff[k + 1]], dataset.read_signal_raw(p))


@pytest.mark.gpu
def test_gpu_rejects_what_the_grammar_excludes(lib):
    for text, code in ((b"12 1_000 5", -1), (b"1 2 3.5", -1), (b"70000", -1), (b"1 - 5", -1), (b"1234567", -1)):
        rc, _, _ = _parse(lib, text, [0, len(text)])
        assert rc == code, (text, rc, lib.ntb_last_error())
    rc, _, _ = _parse(lib, b"1 2 3 4 5 6 7 8", [0, 15], capacity=3)
    assert rc == -4
    rc, samples, off = _parse(lib, b"", [0, 0])
    assert rc == 0 and len(samples) == 0
    rc, samples, off = _parse(lib, b"7 8", [0, 1, 1, 3])          # a token never spans two files: "7", "", "8"
    assert rc == 0 and list(samples) == [7, 8] and list(off) == [0, 1, 1, 2]


@pytest.mark.gpu
def test_file_bytes_to_model_input_on_the_device(lib, tmp_path):
    """Rank 2 chained into row a9 without leaving the GPU: file bytes -> ntb_parse_signal -> ntb_pack_windows (z-score +
    windows) against the host route FullSignalSequence.load_file_windows (read_signal + windows): window lengths
    identical, samples within the float32 z-score tolerance of the packer's own parity test (2e-6)."""
    import torch
    rng = np.random.default_rng(12)
    T, step = 100, 100
    paths, blobs = [], []
    for f in range(5):
        vals = rng.integers(350, 800, int(rng.integers(150, 4000)))
        p = tmp_path / ("r%d.signal" % f)
        p.write_text(" ".join(map(str, vals)) + "\n")
        paths.append(str(p))
        blobs.append(p.read_bytes())
    file_off = np.concatenate([[0], np.cumsum([len(b) for b in blobs])]).astype(np.int64)
    total = int(file_off[-1])
    text = torch.as_tensor(np.frombuffer(b"".join(blobs), dtype=np.uint8).copy()).cuda()
    cap = total // 2 + len(paths) + 1
    ws = torch.empty(lib.ntb_parse_signal_workspace_bytes(total, len(paths)), dtype=torch.uint8, device="cuda")
    samples = torch.empty(cap, dtype=torch.int16, device="cuda")
    read_off = torch.empty(len(paths) + 1, dtype=torch.int64, device="cuda")
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(lib.ntb_parse_signal(text.data_ptr(), file_off.ctypes.data, len(paths), total, ws.data_ptr(), ws.numel(), cap,
                                    samples.data_ptr(), read_off.data_ptr(), status.data_ptr(), st))
    ro = read_off.cpu().numpy()                                   # the host sizes the window table (as for any job)
    assert int(status[0]) == 0
    counts = [-(-int(n) // step) for n in np.diff(ro)]
    win_off = torch.as_tensor(np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)).cuda()
    n_win = int(sum(counts))
    x = torch.empty((n_win, T), dtype=torch.float32, device="cuda")
    ln = torch.empty(n_win, dtype=torch.int32, device="cuda")
    _lib.check(lib.ntb_pack_windows(samples.data_ptr(), read_off.data_ptr(), win_off.data_ptr(), len(paths), n_win, T, step, 0,
                                    x.data_ptr(), ln.data_ptr(), st))
    torch.cuda.synchronize()
    ds = dataset.FullSignalSequence(paths, mode=2, batch_size=10, seq_len=T, step=step, verbose=False)
    xs, ls = zip(*(ds.load_file_windows(p) for p in paths))
    assert np.array_equal(ln.cpu().numpy(), np.concatenate(ls))
    np.testing.assert_allclose(x.cpu().numpy(), np.concatenate(xs), atol=2e-6, rtol=0)
