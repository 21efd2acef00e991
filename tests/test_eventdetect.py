"""Event detection (second half; nanonet eventdetection/filters.py::minknow_event_detect).

Parity PINNED twice over: tests/golden/event_detect.npz holds the outputs of the reference's own Python + C
(tests/golden/make_golden_eventdetect.py; filters.c compiled unmodified by oracle/build_ref.sh), and where
oracle/_ref/libnanonetfilters.so is present (it travels with the repo snapshot) its C functions are also run live on
fresh inputs.  Bar: event boundaries (start, length) and means bit-identical; stdv bit-identical except where the
reference's ``ev_mean ** 2`` (libm pow) and a plain multiply differ in the last bit of the variance -- allowed 4 ulp of
the mean square, relative to the variance, and counted.
"""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_eventdetect as OD
from nanopore_rnn_b200 import _lib, eventdetection

from conftest import GOLDEN, ROOT

REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libnanonetfilters.so")


@pytest.fixture(scope="module")
def golden():
    g = np.load(os.path.join(GOLDEN, "event_detect.npz"))
    return {k: g[k] for k in g.files}


def params(g, p):
    par = g["params_%d" % p]
    return [int(par[0]), int(par[1])], [float(par[2]), float(par[3])], float(par[4])


def squiggle(rng, n, dwell=9.0, noise=1.2):
    out = np.empty(n)
    i = 0
    while i < n:
        d = 1 + int(rng.geometric(1.0 / dwell))
        out[i:i + d] = rng.uniform(55.0, 125.0)
        i += d
    return out + rng.normal(0.0, noise, n)


def check_events(got, want):
    """start / length / mean identical; stdv identical up to the pow-vs-multiply last bit (see module docstring)."""
    assert got.shape == want.shape
    assert np.array_equal(got[:, :3], want[:, :3])
    diff = np.abs(got[:, 3] ** 2 - want[:, 3] ** 2)
    assert np.all(diff <= 4 * np.spacing(want[:, 2] ** 2) + 4 * np.spacing(want[:, 3] ** 2))
    return int((got[:, 3] != want[:, 3]).sum())


# ---------------------------------------------------------------- CPU
@pytest.mark.parametrize("p", [0, 1])
def test_oracle_matches_reference_golden_bit_for_bit(golden, p):
    wl, th, ph = params(golden, p)
    off, eo, po = golden["read_off"], golden["event_off_%d" % p], golden["peak_off_%d" % p]
    for r in range(len(off) - 1):
        ev, pk = OD.minknow_event_detect(golden["raw"][off[r]:off[r + 1]], float(golden["sample_rate"]), wl, th, ph)
        assert np.array_equal(ev, golden["events_%d" % p][eo[r]:eo[r + 1]])
        assert np.array_equal(pk, golden["peaks_%d" % p][po[r]:po[r + 1]])


@pytest.mark.skipif(not os.path.exists(REF_LIB), reason="oracle/_ref not built (oracle/build_ref.sh needs /root/reference)")
def test_oracle_matches_compiled_reference_c_live():
    """compute_sum_sumsq and compute_tstat of the reference's own C against the restatement, on fresh data."""
    ref = C.CDLL(REF_LIB)
    x = squiggle(np.random.default_rng(1), 30000)
    s, q = np.empty_like(x), np.empty_like(x)
    dp = C.POINTER(C.c_double)
    ref.compute_sum_sumsq(x.ctypes.data_as(dp), s.ctypes.data_as(dp), q.ctypes.data_as(dp), C.c_size_t(len(x)))
    so, qo = OD.compute_sum_sumsq(x)
    assert np.array_equal(s, so) and np.array_equal(q, qo)
    for w in (3, 6, 16, 40):
        t = np.empty_like(x)
        ref.compute_tstat(s.ctypes.data_as(dp), q.ctypes.data_as(dp), t.ctypes.data_as(dp), C.c_size_t(len(x)),
                          C.c_size_t(w), C.c_bool(False))
        assert np.array_equal(t, OD.compute_tstat(so, qo, w))


def test_unsupported_options_raise():
    with pytest.raises(NotImplementedError):
        eventdetection.minknow_event_detect(np.ones(100), 4000.0, get_peaks=True)
    with pytest.raises(NotImplementedError):
        eventdetection.minknow_event_detect(np.ones(100), None)


# ---------------------------------------------------------------- GPU
@pytest.fixture(scope="module")
def lib():
    lib = _lib.load()
    assert lib.ntb_device_ok() == 1, lib.ntb_last_error()
    return lib


@pytest.mark.gpu
@pytest.mark.parametrize("p", [0, 1])
def test_gpu_matches_reference_golden(lib, golden, p):
    wl, th, ph = params(golden, p)
    off, eo = golden["read_off"], golden["event_off_%d" % p]
    raws = [golden["raw"][a:b] for a, b in zip(off[:-1], off[1:])]
    evs = eventdetection.minknow_event_detect_batch(raws, float(golden["sample_rate"]), wl, th, ph)
    assert [len(e) for e in evs] == list(np.diff(eo))
    last_bit = 0
    for r, ev in enumerate(evs):
        got = np.stack([ev["start"], ev["length"], ev["mean"], ev["stdv"]], axis=1)
        last_bit += check_events(got, golden["events_%d" % p][eo[r]:eo[r + 1]])
    assert last_bit <= 0.001 * eo[-1] + 1, last_bit           # pow vs multiply: rare
    # the reference's argument order does not matter: the shorter window dominates either way
    ev_rev = eventdetection.minknow_event_detect(raws[0], float(golden["sample_rate"]), wl[::-1], th[::-1], ph)
    assert np.array_equal(ev_rev["mean"], evs[0]["mean"])


@pytest.mark.gpu
@pytest.mark.parametrize("lanes", ["0", "1", "4", "8", "16", "4:fused", "8:fused"])
def test_both_kernel_forms_agree_with_the_golden(golden, lanes):
    """The warp-per-read and the R-reads-per-warp forms (R = 4, 8, 16, 32) of the sequential stages (the library picks by read count; the switch is
    read once per process, hence the subprocess)."""
    import subprocess
    import sys
    code = ("import sys, numpy as np; sys.path[:0] = [%r, %r]; import nanopore_rnn_b200; "
            "from nanopore_rnn_b200 import eventdetection as E; g = np.load(%r); o = g['read_off']; eo = g['event_off_0']; "
            "ev = E.minknow_event_detect_batch([g['raw'][a:b] for a, b in zip(o[:-1], o[1:])], 4000.0, [3, 6], [1.4, 1.1], 0.2); "
            "m = np.concatenate([e['mean'] for e in ev]); s = np.concatenate([e['start'] for e in ev]); "
            "assert np.array_equal(m, g['events_0'][:, 2]) and np.array_equal(s, g['events_0'][:, 0]); print('ok')"
            % (ROOT, os.path.join(ROOT, "oracle"), os.path.join(GOLDEN, "event_detect.npz")))
    env = dict(os.environ, NTB_ED_LANES=lanes.split(":")[0])
    if lanes.endswith(":fused"):                    # the opt-in one-kernel form of cumulative sums + t-statistics + peaks
        env["NTB_ED_FUSED"] = "1"
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True,
                         text=True, timeout=300)
    assert out.returncode == 0 and "ok" in out.stdout, out.stderr[-2000:]


RAGGED_CODE = """
import sys, numpy as np
sys.path[:0] = [%r, %r]
import nanopore_rnn_b200
from nanopore_rnn_b200 import eventdetection as E
rng = np.random.default_rng(19)
def squiggle(n, dwell=9.0, noise=1.2):
    out = np.empty(n); i = 0
    while i < n:
        d = 1 + int(rng.geometric(1.0 / dwell)); out[i:i + d] = rng.uniform(55.0, 125.0); i += d
    return out + rng.normal(0.0, noise, n)
raws = [squiggle(int(n)) for n in rng.integers(1, 3000, 257)] + [squiggle(70000, dwell=12.0), np.full(500, 80.0), squiggle(33)]
res = {}
for tag, wl, th in (("a", [3, 6], [1.4, 1.1]), ("b", [5, 31], [1.2, 1.0]), ("c", [6, 3], [1.1, 1.4])):
    evs = E.minknow_event_detect_batch(raws, 4000.0, wl, th, 0.2)
    res["n_" + tag] = np.array([len(e) for e in evs])
    for k in ("start", "length", "mean", "stdv"):
        res[k + "_" + tag] = np.concatenate([e[k] for e in evs])
np.savez(sys.argv[1], **res)
print("ok")
"""


@pytest.mark.gpu
def test_fused_kernel_equals_the_separate_kernels_on_ragged_reads(tmp_path):
    """k_ed_fused (opt-in: cumulative sums + t-statistics + peaks in one kernel, 4 / 8 reads per warp) against the warp-per-read
    chain of separate kernels: 260 ragged reads (1 ... 70 000 samples, a flat one, one shorter than the windows), three
    parameter sets incl. the largest window the fused form takes (31) and the reversed argument order.  Bit-identical."""
    import subprocess
    import sys
    outs = {}
    for tag, env in (("sep", {"NTB_ED_LANES": "0"}), ("f4", {"NTB_ED_LANES": "4", "NTB_ED_FUSED": "1"}),
                     ("f8", {"NTB_ED_LANES": "8", "NTB_ED_FUSED": "1"}), ("u8", {"NTB_ED_LANES": "8"})):
        path = str(tmp_path / (tag + ".npz"))
        r = subprocess.run([sys.executable, "-c", RAGGED_CODE % (ROOT, os.path.join(ROOT, "oracle")), path],
                           env=dict(os.environ, **env), capture_output=True, text=True, timeout=600)
        assert r.returncode == 0 and "ok" in r.stdout, r.stderr[-2000:]
        outs[tag] = np.load(path)
    for tag in ("f4", "f8", "u8"):
        for k in outs["sep"].files:
            assert np.array_equal(outs[tag][k], outs["sep"][k]), (tag, k)
    assert int(outs["sep"]["n_a"].sum()) > 5000


@pytest.mark.gpu
def test_gpu_matches_oracle_on_many_ragged_reads(lib):
    rng = np.random.default_rng(9)
    raws = [squiggle(rng, int(n)) for n in rng.integers(1, 6000, 300)] + [squiggle(rng, 150000, dwell=12.0)]
    evs = eventdetection.minknow_event_detect_batch(raws, 4000.0, [3, 6], [1.4, 1.1], 0.2)
    for i in list(range(0, 300, 7)) + [300]:
        want, _ = OD.minknow_event_detect(raws[i], 4000.0, [3, 6], [1.4, 1.1], 0.2)
        got = np.stack([evs[i][k] for k in ("start", "length", "mean", "stdv")], axis=1)
        check_events(got, want)
    # a read's events do not depend on the other reads of the call
    alone = eventdetection.minknow_event_detect(raws[5], 4000.0, [3, 6], [1.4, 1.1], 0.2)
    assert np.array_equal(alone["mean"], evs[5]["mean"]) and np.array_equal(alone["stdv"], evs[5]["stdv"])


@pytest.mark.gpu
def test_raw_to_features_chain_matches_the_reference_chain(lib, golden):
    """Events straight into events_to_features (the two halves of rank 4 back to back): boundaries are identical to the
    reference's, so the features built on the reference's events and on ours agree to the stdv last-bit tolerance."""
    from nanopore_rnn_b200 import features
    import oracle_events as OE
    wl, th, ph = params(golden, 0)
    off = golden["read_off"]
    raw = golden["raw"][off[0]:off[1]]
    ev = eventdetection.minknow_event_detect(raw, 4000.0, wl, th, ph)
    X = features.events_to_features(ev)
    want_ev, _ = OD.minknow_event_detect(raw, 4000.0, wl, th, ph)
    want = OE.events_to_features(want_ev[:, 2], want_ev[:, 3], want_ev[:, 1])
    assert X.shape == want.shape and np.allclose(X, want, rtol=0, atol=1e-9)
    assert np.array_equal(X[:, [4, 6]], want[:, [4, 6]])    # mean and length columns: bit-identical end to end


@pytest.mark.gpu
def test_device_entry_point_capacity_and_errors(lib):
    import torch
    raw = squiggle(np.random.default_rng(4), 5000)
    off = np.asarray([0, 2000, 5000], dtype=np.int64)
    d = torch.as_tensor(raw).cuda()
    wb = lib.ntb_event_detect_workspace_bytes(2, 5000)
    ws = torch.empty(wb, dtype=torch.uint8, device="cuda")
    ev = torch.zeros((5000, 4), dtype=torch.float64, device="cuda")
    eo = torch.zeros(3, dtype=torch.int64, device="cuda")
    status = torch.full((1,), 9, dtype=torch.int32, device="cuda")
    wl, th = np.asarray([3, 6], dtype=np.int32), np.asarray([1.4, 1.1])

    def run(cap, wbytes=wb, table=off):
        return lib.ntb_event_detect(d.data_ptr(), table.ctypes.data, 2, 5000, 4000.0, wl.ctypes.data, th.ctypes.data, 0.2,
                                    ws.data_ptr(), wbytes, cap, ev.data_ptr(), eo.data_ptr(), status.data_ptr(), None)
    assert run(5000) == 0
    torch.cuda.synchronize()
    assert int(status[0]) == 0
    want0, _ = OD.minknow_event_detect(raw[:2000], 4000.0, [3, 6], [1.4, 1.1], 0.2)
    n0 = int(eo[1])
    assert n0 == len(want0)
    check_events(ev[:n0].cpu().numpy(), want0)
    assert run(10) == 0
    torch.cuda.synchronize()
    assert int(status[0]) == 1                               # more events than slots: flagged, nothing out of bounds
    assert run(5000, 64) == -4
    assert run(5000, wb, np.asarray([0, 3000, 2000], dtype=np.int64)) == -1
