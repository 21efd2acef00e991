"""Event-feature front end (nanonet/nanonet/features.py::events_to_features).

This row's parity is PINNED: tests/golden/event_features.npz holds the outputs of the reference's own functions
(tests/golden/make_golden_events.py executes them), and both the oracle restatement and the CUDA kernel are held to them
BIT FOR BIT in float64 -- the kernel performs NumPy's operations in NumPy's order, so no tolerance is needed; the float32
output must equal the golden rounded once (the reference's ``.astype(nn.dtype)``, nanonetcall.py:204).
"""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_events as OE
from nanopore_rnn_b200 import _lib, features

from conftest import GOLDEN


@pytest.fixture(scope="module")
def golden():
    g = np.load(os.path.join(GOLDEN, "event_features.npz"))
    return {k: g[k] for k in g.files}


def synthetic_read(n, seed, offset=0.0):
    rng = np.random.default_rng(seed)
    return {"mean": rng.choice(np.linspace(60.0, 120.0, 64), n) + rng.normal(0.0, 1.5, n) + offset,
            "stdv": np.abs(rng.normal(1.5, 0.4, n)) + 0.2,
            "length": rng.integers(3, 60, n) / 4000.0}


# ---------------------------------------------------------------- CPU: the oracle against the reference's outputs
def test_oracle_matches_reference_golden_bit_for_bit(golden):
    X = OE.events_to_features_reads(golden["mean"], golden["stdv"], golden["length"], golden["read_off"])
    assert X.dtype == np.float64 and X.shape == golden["features"].shape
    assert np.array_equal(X, golden["features"])


def test_golden_covers_the_edge_cases(golden):
    off, F = golden["read_off"], golden["features"]
    lens = np.diff(off)
    assert 1 in lens and 2 in lens and lens.max() >= 5000
    one = int(np.where(lens == 1)[0][0])
    assert np.array_equal(F[off[one]], np.zeros(12))                   # a single event standardises to all zeros
    const = int(np.where(lens == 257)[0][0])                          # constant stdv field: scale 0 -> 1, column stays 0
    assert np.all(F[off[const]:off[const + 1], 5] == 0.0)
    for r in range(len(lens)):                                        # zero padding past both read ends
        a, b = off[r], off[r + 1]
        assert np.all(F[a, 0:4] == 0.0) and np.all(F[b - 1, 8:12] == 0.0)


@pytest.mark.parametrize("n", [1, 7, 8, 9, 127, 128, 129, 136, 255, 1000, 4099])
def test_pairwise_sum_is_numpys_order(n):
    a = np.random.default_rng(n).normal(0, 1e3, n)
    assert OE.pairwise_sum(a) == np.add.reduce(a)
    assert OE.pairwise_sum(a[::-1]) == np.add.reduce(a[::-1])


def test_unsupported_options_raise():
    ev = synthetic_read(16, 0)
    with pytest.raises(NotImplementedError):
        features.events_to_features(ev, sloika_model=True)
    with pytest.raises(NotImplementedError):
        features.events_to_features(ev, window=[-2, -1, 0, 1, 2])


# ---------------------------------------------------------------- GPU: the kernel through the C ABI
def _run(lib, mean, stdv, length, read_off, f64=True, host=False):
    import torch
    total = int(read_off[-1])
    off = np.ascontiguousarray(read_off, dtype=np.int64)
    out_dt = np.float64 if f64 else np.float32
    if host:
        out = np.empty((total, 12), dtype=out_dt)
        m, s, l = (np.ascontiguousarray(v, dtype=np.float64) for v in (mean, stdv, length))
        _lib.check(lib.ntb_event_features_host(m.ctypes.data, s.ctypes.data, l.ctypes.data, off.ctypes.data,
                                               len(off) - 1, total, int(f64), out.ctypes.data, None))
        return out
    d = [torch.as_tensor(np.ascontiguousarray(v, dtype=np.float64)).to("cuda:0") for v in (mean, stdv, length)]
    out = torch.full((total, 12), float("nan"), dtype=torch.float64 if f64 else torch.float32, device="cuda:0")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    wb = lib.ntb_event_features_workspace_bytes(len(off) - 1, total)
    ws = torch.empty(max(int(wb), 256), dtype=torch.uint8, device="cuda:0")
    _lib.check(lib.ntb_event_features(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), off.ctypes.data,
                                      len(off) - 1, total, int(f64), ws.data_ptr(), wb, out.data_ptr(), st))
    torch.cuda.synchronize()
    return out.cpu().numpy()


@pytest.fixture(scope="module")
def lib():
    lib = _lib.load()
    assert lib.ntb_device_ok() == 1, lib.ntb_last_error()
    return lib


@pytest.mark.gpu
@pytest.mark.parametrize("host", [False, True])
def test_kernel_matches_reference_golden_bit_for_bit(lib, golden, host):
    X = _run(lib, golden["mean"], golden["stdv"], golden["length"], golden["read_off"], f64=True, host=host)
    assert np.array_equal(X, golden["features"])
    X32 = _run(lib, golden["mean"], golden["stdv"], golden["length"], golden["read_off"], f64=False, host=host)
    assert X32.dtype == np.float32 and np.array_equal(X32, golden["features"].astype(np.float32))


@pytest.mark.gpu
def test_kernel_every_read_length_up_to_300(lib):
    """Every shape of the summation tree's first levels (n < 8, the 8-lane blocks with and without a tail, the first
    splits) against the explicit-order oracle, all reads in one launch."""
    reads = [synthetic_read(n, 1000 + n) for n in range(1, 301)]
    off = np.concatenate([[0], np.cumsum([len(r["mean"]) for r in reads])])
    cat = {k: np.concatenate([r[k] for r in reads]) for k in ("mean", "stdv", "length")}
    want = OE.events_to_features_reads(cat["mean"], cat["stdv"], cat["length"], off)
    got = _run(lib, cat["mean"], cat["stdv"], cat["length"], off)
    assert np.array_equal(got, want)


@pytest.mark.gpu
def test_kernel_long_and_ragged_reads(lib, monkeypatch):
    """Reads of 9000 events (the reference's max_len) up to 300 000 -- all three kernel classes in one call (field in
    shared memory at two CTAs / one CTA per SM, any-length path) -- empty reads in between, an offset of 1e6 pA that
    makes the re-centring branches fire; NumPy's own reductions are the checker at these sizes."""
    monkeypatch.setattr(OE, "_SUM", np.add.reduce)
    lens = [9000, 0, 65537, 1, 300000, 0, 12345, 8999, 20000, 27001, 14000, 13000, 28500]
    reads = [synthetic_read(n, 7 + i, offset=1.0e6 if i % 2 else 0.0) for i, n in enumerate(lens)]
    off = np.concatenate([[0], np.cumsum(lens)])
    cat = {k: np.concatenate([r[k] for r in reads]) for k in ("mean", "stdv", "length")}
    want = OE.events_to_features_reads(cat["mean"], cat["stdv"], cat["length"], off)
    got = _run(lib, cat["mean"], cat["stdv"], cat["length"], off)
    assert np.array_equal(got, want)


@pytest.mark.gpu
def test_host_mirror_and_batch_invariance(lib, golden):
    """features.events_to_features (the reference's signature) on a structured event table, and: a read's features
    do not depend on which other reads share the launch."""
    off = golden["read_off"]
    ev = np.zeros(off[1] - off[0], dtype=[("start", "f8"), ("length", "f8"), ("mean", "f8"), ("stdv", "f8")])
    for k in ("mean", "stdv", "length"):
        ev[k] = golden[k][off[0]:off[1]]
    X = features.events_to_features(ev)
    assert X.dtype == np.float64 and np.array_equal(X, golden["features"][off[0]:off[1]])
    reads = [{k: golden[k][a:b] for k in ("mean", "stdv", "length")} for a, b in zip(off[:-1], off[1:])]
    Xb, ob = features.events_to_features_batch(reads[::-1])
    assert np.array_equal(ob, np.concatenate([[0], np.cumsum(np.diff(off)[::-1])]))
    assert np.array_equal(Xb[ob[-2]:ob[-1]], golden["features"][off[0]:off[1]])
    X32, _ = features.events_to_features_batch(reads, dtype=np.float32)
    assert np.array_equal(X32, golden["features"].astype(np.float32))


@pytest.mark.gpu
def test_bad_read_tables_are_rejected(lib):
    import torch
    z = torch.zeros(16, dtype=torch.float64, device="cuda:0")
    out = torch.zeros(16 * 12, dtype=torch.float64, device="cuda:0")
    for off in ([0, 8, 4, 16], [1, 16], [0, 8]):
        o = np.asarray(off, dtype=np.int64)
        rc = lib.ntb_event_features(z.data_ptr(), z.data_ptr(), z.data_ptr(), o.ctypes.data, len(o) - 1, 16, 1,
                                    out.data_ptr(), out.numel() * 8, out.data_ptr(), None)
        assert rc == -1, (off, rc)
    o = np.asarray([0, 16], dtype=np.int64)                            # a workspace that is too small: NTB_ENOMEM
    rc = lib.ntb_event_features(z.data_ptr(), z.data_ptr(), z.data_ptr(), o.ctypes.data, 1, 16, 1, z.data_ptr(), 64,
                                out.data_ptr(), None)
    assert rc == -4 and b"workspace" in lib.ntb_last_error()
    assert features.events_to_features_batch([])[0].shape == (0, 12)
