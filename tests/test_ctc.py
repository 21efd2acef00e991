"""CTC decoders: the oracle (oracle/oracle_ctc.py, a restatement of TensorFlow 1.4's
CTCGreedyDecoder / CTCBeamSearchDecoder -- parity unpinned, TF is absent) checked independently by exhaustive path
enumeration (CPU), and the CUDA kernels checked against the oracle (-m gpu)."""
import numpy as np
import pytest

import oracle_ctc as O
import oracle_np as A


def _logits(seed, n, T, C, scale=2.0):
    return (scale * np.random.default_rng(seed).standard_normal((n, T, C))).astype(np.float32)


# ------------------------------------------------------------------------------------------ CPU: the oracle itself
def test_greedy_oracle_matches_the_forward_oracle_and_sums_the_maxima():
    lg = _logits(1, 9, 40, 5)
    lens = np.array([40, 0, 1, 17, 40, 39, 2, 40, 8], np.int32)
    lab, llen, neg = O.ctc_greedy(lg, lens)
    ref, ref_neg = A.ctc_greedy(lg, lens)                         # dense int64, -1 padded (network.py:531-532)
    assert np.array_equal(neg, ref_neg)
    assert np.array_equal(lab[:, :ref.shape[1]], ref) and (lab[:, ref.shape[1]:] == -1).all()
    for b in range(len(lg)):
        assert abs(neg[b] + lg[b, :lens[b]].max(axis=1).astype(np.float64).sum()) < 1e-4
    lab2, llen2, _ = O.ctc_greedy(lg, lens, merge_repeated=False)
    assert (llen2 >= llen).all() and llen2.sum() > llen.sum()


@pytest.mark.parametrize("log_softmax", [False, True])
def test_beam_oracle_equals_exhaustive_enumeration_when_nothing_is_pruned(log_softmax):
    """T <= 6, C <= 4: every collapsed label sequence's total must be the log of its summed path products, the top beam
    the arg-max sequence, top_paths the sorted list."""
    for seed, (T, C) in enumerate([(1, 2), (3, 3), (5, 3), (6, 4), (4, 4)]):
        lg = _logits(10 + seed, 1, T, C, scale=1.5)[0]
        truth = O.brute_force_sequences(lg, log_softmax)
        k = min(len(truth), 6)
        seqs, probs = O.beam_search_one(lg, beam_width=10000, top_paths=k, merge_repeated=False, log_softmax=log_softmax)
        want = sorted(truth.items(), key=lambda kv: -kv[1])[:k]
        assert [tuple(s) for s in seqs] == [s for s, _ in want]
        assert np.allclose(probs, [p for _, p in want], atol=2e-5)


def test_beam_oracle_narrow_beam_and_merge_repeated():
    lg = _logits(3, 1, 30, 5)[0]
    s1, p1 = O.beam_search_one(lg, beam_width=1, top_paths=1)
    s100, p100 = O.beam_search_one(lg, beam_width=100, top_paths=3)
    assert len(s100) == 3 and p100[0] >= p100[1] >= p100[2] and p100[0] >= p1[0] - 1e-6
    sm, _ = O.beam_search_one(lg, beam_width=100, top_paths=1, merge_repeated=True)
    assert all(a != b for a, b in zip(sm[0], sm[0][1:]))
    # empty sequence: the root alone
    s0, p0 = O.beam_search_one(lg[:0], beam_width=100, top_paths=2)
    assert s0 == [[]] and p0 == [0.0]


# ------------------------------------------------------------------------------------------ GPU: kernels vs the oracle
@pytest.mark.gpu
def test_ctc_greedy_kernel_matches_oracle_bit_for_bit():
    from nanopore_rnn_b200 import ctc
    for C, T, n in [(4, 100, 70), (6, 33, 5), (2, 300, 40), (8, 64, 33)]:
        lg = _logits(C * 7 + T, n, T, C)
        lg[0, :, :] = lg[0, 0, 0]                                  # exact ties everywhere: first max wins
        lens = np.random.default_rng(T).integers(0, T + 1, n).astype(np.int32)
        lens[:2] = [T, 0]
        for merge in (True, False):
            dec, neg = ctc.ctc_greedy_decoder(lg, lens, merge_repeated=merge, time_major=False)
            lab, llen, rneg = O.ctc_greedy(lg, lens, merge)
            w = int(llen.max())
            assert dec[0].shape == (n, w) and np.array_equal(dec[0], lab[:, :w])
            assert np.array_equal(neg[:, 0], rneg)                 # same float32 additions in the same order
        # time-major input as TensorFlow takes it
        dec_tm, _ = ctc.ctc_greedy_decoder(np.transpose(lg, (1, 0, 2)), lens)
        assert np.array_equal(dec_tm[0], O.ctc_greedy(lg, lens)[0][:, :dec_tm[0].shape[1]])


@pytest.mark.gpu
@pytest.mark.parametrize("C,T,n,W,k,log_softmax", [(5, 40, 24, 100, 3, False), (4, 100, 12, 100, 1, False), (6, 25, 16, 4, 2, False),
                                                   (3, 12, 30, 128, 5, True), (8, 20, 9, 10, 4, False), (5, 60, 8, 1, 1, False)])
def test_ctc_beam_search_kernel_matches_oracle(C, T, n, W, k, log_softmax):
    from nanopore_rnn_b200 import ctc
    lg = _logits(100 + C + T, n, T, C)
    lens = np.random.default_rng(T + C).integers(1, T + 1, n).astype(np.int32)
    lens[:2] = [T, 0]
    for merge in (False, True):
        dec, logp = ctc.ctc_beam_search_decoder(lg, lens, beam_width=W, top_paths=k, merge_repeated=merge,
                                                time_major=False, log_softmax=log_softmax)
        rp, rl, rlp = O.ctc_beam_search(lg, lens, W, k, merge, log_softmax)
        assert logp.shape == (n, k) and len(dec) == k
        for b in range(n):
            for j in range(k):
                if not np.isfinite(rlp[b, j]):
                    assert not np.isfinite(logp[b, j])
                    continue
                got = [int(v) for v in dec[j][b] if v >= 0]
                want = [int(v) for v in rp[b, j, :rl[b, j]]]
                # float32 log1p / exp of the device library differ from the host's in the last place: a swap of two beams is
                # accepted only where the oracle's own totals are that close
                near = [abs(rlp[b, j] - rlp[b, i]) < 2e-4 for i in range(k) if i != j and np.isfinite(rlp[b, i])]
                assert abs(logp[b, j] - rlp[b, j]) < 2e-4 * max(1.0, abs(rlp[b, j])), (b, j, logp[b, j], rlp[b, j])
                assert got == want or any(near), (b, j, got, want)


@pytest.mark.gpu
def test_ctc_beam_search_kernel_finds_the_enumerated_optimum():
    from nanopore_rnn_b200 import ctc
    lg = _logits(77, 20, 6, 4, scale=1.5)
    lens = np.full(20, 6, np.int32)
    dec, logp = ctc.ctc_beam_search_decoder(lg, lens, beam_width=128, top_paths=1, merge_repeated=False, time_major=False)
    for b in range(20):
        truth = O.brute_force_sequences(lg[b])
        best, p = max(truth.items(), key=lambda kv: kv[1])
        second = sorted(truth.values())[-2]
        got = tuple(int(v) for v in dec[0][b] if v >= 0)
        assert abs(logp[b, 0] - p) < 1e-4
        assert got == best or p - second < 1e-4
