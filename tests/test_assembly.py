"""Consensus assembly (nanotensor/dataset.py:421-432 process_output).

This row's parity is PINNED: tests/golden/assembly.npz holds the outputs of the reference's own index2base /
simple_assembly (executed verbatim by tests/golden/make_golden_assembly.py).  Integer work: the oracle restatement and
the CUDA path must reproduce the consensus AND the vote matrix exactly.
"""
import ctypes as C
import difflib
import os

import numpy as np
import pytest

import oracle_assembly as OA
from nanopore_rnn_b200 import _lib, assembly, dataset

from conftest import GOLDEN


@pytest.fixture(scope="module")
def golden():
    g = np.load(os.path.join(GOLDEN, "assembly.npz"))
    return {k: g[k] for k in g.files}


def sliding_job(rng, n_reads, n_windows, T=100, err=0.05):
    """Label rows of windows sliding over random sequences, with errors; returns labels, read_off."""
    rows, off = [], [0]
    for _ in range(n_reads):
        nw = int(rng.integers(1, n_windows + 1))
        truth = rng.integers(0, 4, 8 * nw + 40)
        for w in range(nw):
            seg = truth[8 * w + rng.integers(0, 3): 8 * w + 24 + rng.integers(0, 6)].copy()
            flip = rng.random(len(seg)) < err
            seg[flip] = rng.integers(0, 5, int(flip.sum()))
            seg = seg[rng.random(len(seg)) >= err / 2]
            row = np.full(T, -1, dtype=np.int32)
            row[:len(seg)] = seg
            rows.append(row)
        off.append(off[-1] + nw)
    return np.asarray(rows), np.asarray(off, dtype=np.int64)


# ---------------------------------------------------------------- CPU
def test_oracle_matches_reference_golden_exactly(golden):
    cons, off, votes = OA.assemble(golden["labels"], golden["read_off"])
    assert np.array_equal(off, golden["cons_off"])
    assert np.array_equal(cons, golden["consensus"])
    assert np.array_equal(votes, golden["counts"])


def test_golden_covers_the_edge_cases(golden):
    lens = np.diff(golden["cons_off"])
    nwin = np.diff(golden["read_off"])
    assert lens[np.where(nwin == 1)[0][0]] == 0            # a one-window read has an EMPTY consensus in the reference
    assert lens.max() > 1000                               # the reference's census padding branch
    assert (golden["label_len"] == 0).any() and (golden["label_len"] == 100).any()


def test_longest_block_is_difflibs_choice():
    rng = np.random.default_rng(5)
    for _ in range(1500):
        la, lb = rng.integers(0, 70, 2)
        alpha = int(rng.integers(1, 6))
        a, b = rng.integers(0, alpha, la).astype(np.int8), rng.integers(0, alpha, lb).astype(np.int8)
        sa, sb = ("".join("ACGTE"[x] for x in s) for s in (a, b))
        m = max(difflib.SequenceMatcher(None, sa, sb).get_matching_blocks(), key=lambda x: x[2])
        assert tuple(m) == OA.longest_block(a, b)
    # 200 elements in the second sequence: autojunk -- the only symbol is popular, no anchor, but difflib still extends the
    # empty match at (0, 0) over equal elements
    sa, sb = "A" * 10, "A" * 200
    m = max(difflib.SequenceMatcher(None, sa, sb).get_matching_blocks(), key=lambda x: x[2])
    assert tuple(m) == OA.longest_block(np.zeros(10, np.int8), np.zeros(200, np.int8)) == (0, 0, 10)


# ---------------------------------------------------------------- GPU
@pytest.fixture(scope="module")
def lib():
    lib = _lib.load()
    assert lib.ntb_device_ok() == 1, lib.ntb_last_error()
    return lib


@pytest.mark.gpu
def test_gpu_matches_reference_golden_exactly(lib, golden):
    seqs, counts = assembly.assemble_reads(golden["labels"], golden["read_off"], want_counts=True)
    off = golden["cons_off"]
    text = golden["consensus"].tobytes().decode()
    assert seqs == [text[a:b] for a, b in zip(off[:-1], off[1:])]
    assert np.array_equal(np.concatenate(counts, axis=1), golden["counts"])


@pytest.mark.gpu
def test_gpu_matches_difflib_host_mirror_and_oracle_on_a_larger_job(lib):
    """20k windows in 64 reads: against the host mirror (dataset.assemble_read, which calls difflib itself) and the
    difflib-free oracle; and a read's consensus does not depend on the other reads of the call."""
    labels, off = sliding_job(np.random.default_rng(11), 64, 640)
    seqs = assembly.assemble_reads(labels, off)
    want = [dataset.assemble_read([labels[a:b]]) for a, b in zip(off[:-1], off[1:])]
    assert seqs == want
    cons, coff, _ = OA.assemble(labels[:off[8]], off[:9])
    assert "".join(seqs[:8]) == cons.tobytes().decode()
    assert assembly.assemble_reads(labels[off[5]:off[6]], np.asarray([0, off[6] - off[5]])) == [seqs[5]]


@pytest.mark.gpu
def test_gpu_error_free_windows_rebuild_the_sequence(lib):
    """Size-independent property at full width: windows cut from one sequence without errors, every overlap longer than
    any repeat, reassemble to exactly the covered part of the sequence."""
    rng = np.random.default_rng(3)
    truth = rng.integers(0, 4, 30 * 4000 + 100)
    rows = np.full((4000, 100), -1, dtype=np.int32)
    for w in range(4000):
        rows[w, :90] = truth[30 * w: 30 * w + 90]
    seq = assembly.assemble_reads(rows, np.asarray([0, 4000]))[0]
    assert seq == "".join("ACGT"[x] for x in truth[:30 * 3999 + 90])


def test_oracle_autojunk_regime_matches_reference_golden_and_difflib(golden_dir):
    """Windows of 200 and more called bases (T = 300): difflib's autojunk regime, pinned to the reference's own functions
    (tests/golden/assembly_long.npz) and to difflib on random strings."""
    import difflib
    g = np.load(os.path.join(golden_dir, "assembly_long.npz"))
    assert (g["label_len"] >= 200).sum() >= 20
    cons, off, votes = OA.assemble(g["labels"], g["read_off"])
    assert np.array_equal(off, g["cons_off"]) and np.array_equal(cons, g["consensus"]) and np.array_equal(votes, g["counts"])
    rng = np.random.default_rng(0)
    for trial in range(120):
        la, lb, k = int(rng.integers(1, 300)), int(rng.integers(200, 300)), int(rng.integers(2, 6))
        a, b = rng.integers(0, k, la), rng.integers(0, k, lb)
        if trial % 3 == 0:
            b = np.concatenate([a[int(rng.integers(0, la)):], rng.integers(0, k, lb)])[:lb]
        if trial % 4 == 0:
            b[rng.choice(lb, 3, replace=False)] = 4
        sa, sb = "".join("ACGTE"[v] for v in a), "".join("ACGTE"[v] for v in b)
        m = max(difflib.SequenceMatcher(None, sa, sb).get_matching_blocks(), key=lambda x: x[2])
        assert tuple(m) == tuple(OA.longest_block(a.astype(np.int8), b.astype(np.int8)))


@pytest.mark.gpu
def test_gpu_autojunk_regime_matches_reference_golden_and_oracle(lib, golden_dir):
    g = np.load(os.path.join(golden_dir, "assembly_long.npz"))
    seqs, counts = assembly.assemble_reads(g["labels"], g["read_off"], want_counts=True)
    assert "".join(seqs) == g["consensus"].tobytes().decode()
    assert np.array_equal(np.concatenate(counts, axis=1), g["counts"])
    # random long windows, related and unrelated, rare symbols as the only anchors
    rng = np.random.default_rng(5)
    T, rows, off = 300, [], [0]
    for r in range(12):
        truth = rng.integers(0, 4, 3000)
        truth[rng.choice(3000, 25, replace=False)] = 4
        nw = int(rng.integers(2, 9))
        for w in range(nw):
            n = int(rng.integers(150, 300))
            s0 = w * int(rng.integers(20, 60))
            seg = truth[s0:s0 + n].copy() if r % 3 else rng.integers(0, 5 if r % 2 else 4, n)
            row = np.full(T, -1, np.int32)
            row[:len(seg)] = seg
            rows.append(row)
        off.append(off[-1] + nw)
    labels, off = np.asarray(rows), np.asarray(off, dtype=np.int64)
    got = assembly.assemble_reads(labels, off)
    want, woff, _ = OA.assemble(labels, off)
    assert "".join(got) == want.tobytes().decode() and [len(x) for x in got] == list(np.diff(woff))


@pytest.mark.gpu
def test_device_entry_point_status_and_errors(lib):
    import torch
    labels, off = sliding_job(np.random.default_rng(2), 3, 10)
    n, T = labels.shape
    cap = n * T
    d = torch.as_tensor(labels).cuda()
    ws_b = lib.ntb_assemble_workspace_bytes(n, len(off) - 1, T, cap)
    ws = torch.empty(ws_b, dtype=torch.uint8, device="cuda")
    cons = torch.zeros(cap, dtype=torch.uint8, device="cuda")
    coff = torch.zeros(len(off), dtype=torch.int64, device="cuda")
    status = torch.full((1,), 7, dtype=torch.int32, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def run(lab, capacity, ws_bytes=ws_b):
        return lib.ntb_assemble_reads(lab.data_ptr(), off.ctypes.data, len(off) - 1, n, T, capacity, ws.data_ptr(),
                                      ws_bytes, cons.data_ptr(), coff.data_ptr(), None, status.data_ptr(), st)
    assert run(d, cap) == 0
    torch.cuda.synchronize()
    assert int(status[0]) == 0
    want, woff, _ = OA.assemble(labels, off)
    assert np.array_equal(coff.cpu().numpy(), woff) and np.array_equal(cons[:woff[-1]].cpu().numpy(), want)
    assert run(d, 16) == 0                                  # fewer slots than called bases: flagged, nothing out of bounds
    torch.cuda.synchronize()
    assert int(status[0]) == 1
    assert run(d, cap, 64) == -4 and b"workspace" in lib.ntb_last_error()
    long_rows = np.zeros((n, 256), dtype=np.int32)          # 256 called bases per window: difflib's autojunk regime is
    seqs = assembly.assemble_reads(long_rows, off)          # answered like the reference answers it (no rejection)
    assert len(seqs) == len(off) - 1 and set("".join(seqs)) <= {"A"}
    assert assembly.assemble_reads(np.zeros((0, 100), np.int32), np.asarray([0])) == []
    assert assembly.assemble_reads(np.full((2, 100), -1, np.int32), np.asarray([0, 1, 2])) == ["", ""]
