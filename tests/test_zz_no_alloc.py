"""ntb_ctc_beam_search_noalloc (include/nanotensor_b200.h): the allocation-free variant of ntb_ctc_beam_search, scratch in a
static device array inside the library.  No reference counterpart.  Checked
through the C ABI against the existing entry point on the same inputs, two calls in a row (the second call finds the first
call's contents in the shared scratch), plus the size limit.  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest
import torch

from nanopore_rnn_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
SCRATCH_BYTES = 8 * 1024 * 1024                       # NTB_NOALLOC_SCRATCH_BYTES
MAX_T_TIMES_BEAM = SCRATCH_BYTES // (4 * 12 * 4) - 1  # NTB_NOALLOC_MAX_T_TIMES_BEAM


def P(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _stream():
    return C.c_void_p(torch.cuda.current_stream(torch.device(DEV)).cuda_stream)


def _case(seed, n, T, Cc):
    rng = np.random.default_rng(seed)
    lg = (2.0 * rng.standard_normal((n, T, Cc))).astype(np.float32)
    lens = rng.integers(1, T + 1, n).astype(np.int32)
    lens[:2] = [T, 0]
    return torch.from_numpy(lg).to(DEV), torch.from_numpy(lens).to(DEV)


def _outputs(n, T, k):
    return (torch.full((n, k, T), -7, dtype=torch.int32, device=DEV), torch.full((n, k), -7, dtype=torch.int32, device=DEV),
            torch.full((n, k), 7.0, dtype=torch.float32, device=DEV))


def _reference(lib, lg, lens, n, T, Cc, W, k, merge, lsm):
    ws = torch.empty(int(lib.ntb_ctc_beam_workspace_bytes(n, T, W)), dtype=torch.uint8, device=DEV)
    out = _outputs(n, T, k)
    _lib.check(lib.ntb_ctc_beam_search(P(lg), P(lens), n, T, Cc, W, k, merge, lsm, P(ws), ws.numel(), P(out[0]), P(out[1]),
                                       P(out[2]), _stream()))
    torch.cuda.synchronize()
    return [o.cpu().numpy() for o in out]


def _noalloc(lib, lg, lens, n, T, Cc, W, k, merge, lsm):
    out = _outputs(n, T, k)                                     # outputs are the caller's; the call itself allocates nothing
    _lib.check(lib.ntb_ctc_beam_search_noalloc(P(lg), P(lens), n, T, Cc, W, k, merge, lsm, P(out[0]), P(out[1]), P(out[2]),
                                               _stream()))
    torch.cuda.synchronize()
    return [o.cpu().numpy() for o in out]


def _same(a, b):
    # same kernel, same per-sequence arithmetic (a sequence's result does not depend on which warp of the grid takes it):
    # paths and lengths identical, log-probabilities identical bit for bit (-inf for missing paths included)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    assert np.array_equal(a[2], b[2])


def test_noalloc_beam_search_matches_workspace_entry_twice_in_a_row():
    lib = _lib.load()
    # call 1: 300 sequences (more than the static scratch holds in flight at T x W = 3000 -> the grid strides), 3 paths
    n, T, Cc, W, k = 300, 60, 5, 50, 3
    lg, lens = _case(1, n, T, Cc)
    want1 = _reference(lib, lg, lens, n, T, Cc, W, k, 0, 0)
    got1 = _noalloc(lib, lg, lens, n, T, Cc, W, k, 0, 0)
    _same(got1, want1)
    # call 2, straight after, other shape and flags: the scratch still holds call 1's trees
    n2, T2, C2, W2, k2 = 37, 100, 4, 100, 1
    lg2, lens2 = _case(2, n2, T2, C2)
    want2 = _reference(lib, lg2, lens2, n2, T2, C2, W2, k2, 1, 1)
    got2 = _noalloc(lib, lg2, lens2, n2, T2, C2, W2, k2, 1, 1)
    _same(got2, want2)
    # and call 1 again: unchanged
    _same(_noalloc(lib, lg, lens, n, T, Cc, W, k, 0, 0), want1)


def test_noalloc_beam_search_rejects_sizes_beyond_the_static_scratch():
    lib = _lib.load()
    T, W = 512, 128
    assert T * W > MAX_T_TIMES_BEAM
    lg, lens = _case(3, 2, T, 4)
    out = _outputs(2, T, 1)
    rc = lib.ntb_ctc_beam_search_noalloc(P(lg), P(lens), 2, T, 4, W, 1, 0, 0, P(out[0]), P(out[1]), P(out[2]), _stream())
    assert rc != 0
    msg = lib.ntb_last_error().decode()
    assert "static scratch" in msg and str(MAX_T_TIMES_BEAM) in msg
    torch.cuda.synchronize()
    assert int(out[1].min()) == -7                               # nothing ran
    # the largest shape under the limit still runs (T = 341, beam 128 -> 43 648 <= 43 689)
    T, W = MAX_T_TIMES_BEAM // 128, 128
    lg, lens = _case(4, 3, T, 4)
    _same(_noalloc(lib, lg, lens, 3, T, 4, W, 2, 0, 0), _reference(lib, lg, lens, 3, T, 4, W, 2, 0, 0))
    # empty job: nothing to do, no error
    assert lib.ntb_ctc_beam_search_noalloc(P(lg), P(lens), 0, T, 4, W, 1, 0, 0, P(out[0]), P(out[1]), P(out[2]), _stream()) == 0
