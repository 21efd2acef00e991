"""TEST INFRASTRUCTURE (oracle) -- CPU restatement of the two CTC decoders the reference calls:

* ``tf.nn.ctc_greedy_decoder(logits, seq_len, merge_repeated=True)`` with BOTH outputs, the decoded labels and
  ``neg_sum_logits`` (nanotensor/network.py:528-533 -- the reference keeps ``[0]`` and drops ``neg_sum_logits``;
  chiron/chiron/chiron_rcnn_train.py:68);
* ``tf.nn.ctc_beam_search_decoder(logits, seq_len, merge_repeated=False, top_paths=k)`` (beam_width 100, TF's default;
  chiron/chiron/chiron_rcnn_train.py:66-67, chiron_cnn_train.py:32, train_test.py:43,83).

The arithmetic lives in TensorFlow 1.4 (``tensorflow/core/util/ctc/ctc_decoder.h::CTCGreedyDecoder`` and
``ctc_beam_search.h::CTCBeamSearchDecoder`` with the default ``BaseBeamScorer``), a dependency that is not in
/root/reference and not installed: **parity unpinned**.  What is restated, from the published source of that version:

greedy: per frame t < seq_len the first-max class; ``neg_sum_logits -= max`` in frame order (float32); a class is emitted
iff it is not the blank (last class) and (not merge_repeated or it differs from the previous frame's class).

beam search, ``Step(raw_input)``: ``input = raw_input - max(raw_input)`` is used as log-probability AS IS (TF 1.4 does
not normalise; ``log_softmax=True`` gives the later TF behaviour); beam entries form a prefix tree, each with
(blank, label, total) log-probabilities; every leaf is updated (label += parent's contribution when the parent is still
in the beam, blank = old total + input[blank]) and pushed back, then every leaf in descending old-total order proposes its
inactive children, which enter while they beat the current bottom of the size-``beam_width`` top list (the evicted bottom
leaves the beam and loses its probabilities).  ``TopPaths`` returns the ``top_paths`` best leaves by total with
``newp.total`` as ``log_probability``; ``merge_repeated`` collapses repeated labels of the OUTPUT sequence.

Independent check (tests/test_ctc.py): exhaustive enumeration of all C^T frame paths on tiny cases -- with a beam wide
enough that nothing is pruned the totals must equal the log of the summed path products per collapsed sequence.
"""
import itertools
import math

import numpy as np

LOG_ZERO = -np.inf


def ctc_greedy(logits, lens, merge_repeated=True):
    """logits [n, T, C] float32, lens [n].  Returns (labels int64 [n, T] padded -1, label_len int32 [n],
    neg_sum_logits float32 [n])."""
    logits = np.asarray(logits, dtype=np.float32)
    n, T, C = logits.shape
    out = np.full((n, T), -1, dtype=np.int64)
    olen = np.zeros(n, dtype=np.int32)
    neg = np.zeros(n, dtype=np.float32)
    for b in range(n):
        prev = -1
        acc = np.float32(0.0)
        k = 0
        for t in range(int(lens[b])):
            row = logits[b, t]
            c = int(np.argmax(row))                      # first max
            acc = np.float32(acc - row[c])
            if c != C - 1 and not (merge_repeated and c == prev):
                out[b, k] = c
                k += 1
            prev = c
        olen[b] = k
        neg[b] = acc
    return out, olen, neg


def _lse(a, b):
    a, b = np.float32(a), np.float32(b)
    if a == LOG_ZERO and b == LOG_ZERO:
        return np.float32(LOG_ZERO)
    if a > b:
        return np.float32(a + np.log1p(np.exp(np.float32(b - a), dtype=np.float32), dtype=np.float32))
    return np.float32(b + np.log1p(np.exp(np.float32(a - b), dtype=np.float32), dtype=np.float32))


class _Entry(object):
    __slots__ = ("parent", "label", "children", "ob", "ol", "ot", "nb", "nl", "nt")

    def __init__(self, parent, label):
        self.parent, self.label, self.children = parent, label, None
        self.ob = self.ol = self.ot = np.float32(LOG_ZERO)
        self.nb = self.nl = self.nt = np.float32(LOG_ZERO)

    def active(self):
        return self.nt != LOG_ZERO

    def reset_new(self):
        self.nb = self.nl = self.nt = np.float32(LOG_ZERO)

    def seq(self, merge_repeated):
        labels, e = [], self
        while e.parent is not None:
            labels.append(e.label)
            e = e.parent
        labels.reverse()
        if not merge_repeated:
            return labels
        out, prev = [], -1
        for c in labels:
            if c != prev:
                out.append(c)
            prev = c
        return out


def _push(leaves, e, width):
    """gtl::TopN::push: keep the `width` best by newp.total (the caller has already reset an evicted bottom)."""
    leaves.append(e)
    if len(leaves) > width:
        worst = min(range(len(leaves)), key=lambda i: (leaves[i].nt, -i))
        leaves.pop(worst)


def beam_search_one(logits, beam_width=100, top_paths=1, merge_repeated=False, log_softmax=False):
    """logits [T_valid, C] of ONE sequence.  Returns (list of label lists, list of log-probabilities)."""
    logits = np.asarray(logits, dtype=np.float32)
    C = logits.shape[1]
    blank = C - 1
    root = _Entry(None, -1)
    root.nb = root.nt = np.float32(0.0)
    leaves = [root]
    for t in range(logits.shape[0]):
        inp = (logits[t] - logits[t].max()).astype(np.float32)
        if log_softmax:
            inp = (inp - np.log(np.exp(inp, dtype=np.float32).sum(dtype=np.float32), dtype=np.float32)).astype(np.float32)
        branches = sorted(leaves, key=lambda e: -e.nt)            # Extract(): descending new probability (stable)
        leaves = []
        for b in branches:
            b.ob, b.ol, b.ot = b.nb, b.nl, b.nt
        for b in branches:
            if b.parent is not None:
                if b.parent.active():
                    prev = b.parent.ob if b.label == b.parent.label else b.parent.ot
                    b.nl = _lse(b.nl, prev)
                b.nl = np.float32(b.nl + inp[b.label])
            b.nb = np.float32(b.ot + inp[blank])
            b.nt = _lse(b.nb, b.nl)
            _push(leaves, b, beam_width)

        def is_candidate(total):
            if not total > LOG_ZERO:
                return False
            return len(leaves) < beam_width or total > min(e.nt for e in leaves)

        for b in branches:
            if not is_candidate(b.ot):
                continue
            if b.children is None:
                b.children = [_Entry(b, c) for c in range(C - 1)]
            for c in b.children:
                if c.active():
                    continue
                c.nb = np.float32(LOG_ZERO)
                prev = b.ob if c.label == b.label else b.ot
                c.nl = np.float32(inp[c.label] + prev)
                c.nt = c.nl
                if is_candidate(c.nt):
                    if len(leaves) == beam_width:
                        worst = min(range(len(leaves)), key=lambda i: (leaves[i].nt, -i))
                        leaves[worst].reset_new()                    # bottom leaves the beam
                        leaves.pop(worst)
                    leaves.append(c)
                else:
                    c.ob = c.ol = c.ot = np.float32(LOG_ZERO)
                    c.reset_new()
    best = sorted(leaves, key=lambda e: -e.nt)[:top_paths]
    return [e.seq(merge_repeated) for e in best], [float(e.nt) for e in best]


def ctc_beam_search(logits, lens, beam_width=100, top_paths=1, merge_repeated=False, log_softmax=False):
    """Batch form: logits [n, T, C], lens [n].  Returns (paths int64 [n, top_paths, T] padded -1, path_len int32
    [n, top_paths], log_prob float32 [n, top_paths]); missing paths (fewer leaves than top_paths) have length 0 and
    log-probability -inf."""
    logits = np.asarray(logits, dtype=np.float32)
    n, T, C = logits.shape
    paths = np.full((n, top_paths, T), -1, dtype=np.int64)
    plen = np.zeros((n, top_paths), dtype=np.int32)
    lp = np.full((n, top_paths), -np.inf, dtype=np.float32)
    for b in range(n):
        seqs, probs = beam_search_one(logits[b, :int(lens[b])], beam_width, top_paths, merge_repeated, log_softmax)
        for k, (s, p) in enumerate(zip(seqs, probs)):
            paths[b, k, :len(s)] = s
            plen[b, k] = len(s)
            lp[b, k] = p
    return paths, plen, lp


def brute_force_sequences(logits, log_softmax=False):
    """All C^T frame paths of ONE short sequence: {collapsed label tuple: log of the summed products of exp(input)} with
    input = logits - max (or log-softmax) per frame, the quantity the beam search accumulates."""
    logits = np.asarray(logits, dtype=np.float64)
    T, C = logits.shape
    inp = logits - logits.max(axis=1, keepdims=True)
    if log_softmax:
        inp = inp - np.log(np.exp(inp).sum(axis=1, keepdims=True))
    acc = {}
    for path in itertools.product(range(C), repeat=T):
        lp = sum(inp[t, c] for t, c in enumerate(path))
        seq, prev = [], -1
        for c in path:
            if c != C - 1 and c != prev:
                seq.append(c)
            prev = c
        key = tuple(seq)
        acc[key] = acc.get(key, 0.0) + math.exp(lp)
    return {k: math.log(v) for k, v in acc.items()}
