"""Graph classes with the reference's names, constructor kwargs and attributes
(nanotensor/network.py:29-550 ``BuildGraph`` / ``CrossEntropy`` / ``CtcLoss``), executing on the
B200 engine instead of building a TensorFlow graph.

Where the reference exposes graph tensors (``.prediction``, ``.output_layer``, ``.accuracy``,
``.batch_size``, ``.sequence_length``, ``.x``) this exposes ``Fetch`` handles; ``graph.run(fetches)``
stands in for ``sess.run(fetches)``: it pulls the next batch from the dataset iterator (raising
``OutOfRangeError`` at the end, as tf does) and evaluates the fetches on the GPU.  One reference
batch is one BN group, so results do not depend on how many batches the engine processes per
launch: ``run_job`` evaluates a whole read set in large BN-group-aligned chunks.
"""
import logging as log

import numpy as np

from . import checkpoint as ckpt
from .dataset import OutOfRangeError  # noqa: F401  (re-exported for drivers)
from .engine import Model, ctc_dense
from .error import Usage


class Fetch(object):
    """Handle naming one output of the graph (stands in for a tf.Tensor)."""

    def __init__(self, graph, name):
        self.graph, self.name = graph, name

    def __repr__(self):
        return "<Fetch %s of %s>" % (self.name, type(self.graph).__name__)


class BuildGraph(object):
    """network.py:29-118.  Subclasses fix the output layer / prediction function."""

    graph_name = "CrossEntropy"

    def __init__(self, x_iterator=None, y_iterator=None, seq_iterator=None, network=None, len_x=1, len_y=1,
                 learning_rate=0, seq_len=0, forget_bias=5.0, reuse=None, mode=0,
                 y_shape=None, x_shape=None, seq_shape=None, cost_function="cross_entropy",
                 summary_name="Training", dataset=None, weights=None, model_path=None, model_name=None,
                 device="cuda:0"):
        if dataset is None:
            assert (x_iterator is None) != (x_shape is None), "Either x_shape or x_iterator must be specified"
            if mode == 0 or mode == 1:
                assert (y_iterator is None) != (y_shape is None), "Either y_shape or y_iterator must be specified"
            assert (seq_iterator is None) != (seq_shape is None), \
                "Either seq_shape or seq_iterator must be specified"
        assert network is not None, "Must specify network structure. [{'type': 'blstm', 'name': 'blstm_layer1', " \
                                    "'size': 128}, ...] "
        if mode == 0:
            raise NotImplementedError("training graphs (mode 0: loss / optimizer) are outside the forward path")
        for layer in network:
            if layer["type"] not in ("blstm", "residual_layer", "chiron_fnn"):
                raise Usage("layer type %r is not supported by the B200 forward path "
                            "(supported: blstm, residual_layer, chiron_fnn)" % layer["type"])
        self.dataset = dataset
        self.x_iterator, self.y_iterator, self.seq_iterator = x_iterator, y_iterator, seq_iterator
        self.x_shape, self.y_shape, self.seq_shape = x_shape, y_shape, seq_shape
        self.len_x, self.len_y, self.seq_len = len_x, len_y, seq_len
        self.learning_rate = learning_rate
        self.forget_bias = forget_bias      # dead in the reference too: the per-layer "bias" key is used
        self.mode = mode
        self.reuse = reuse
        self.summary_name = summary_name
        self.network = [dict(l) for l in network]
        self.device = device
        self.model_name = model_name
        self._model = None
        self._weights = None
        # graph "tensors"
        self.x = Fetch(self, "x")
        self.sequence_length = Fetch(self, "sequence_length")
        self.batch_size = Fetch(self, "batch_size")
        self.output_layer = Fetch(self, "output_layer")
        self.prediction = Fetch(self, "prediction")
        if self.mode == 1:
            self.y = Fetch(self, "y")
            self.accuracy = Fetch(self, "accuracy")
        log.info("Graph %s: seq_len %d len_x %d classes %d", type(self).__name__, seq_len, len_x, self.len_y)
        if weights is not None:
            self.set_weights(weights)
        elif model_path is not None:
            self.restore(model_path)

    # ------------------------------------------------------------------ weights (Saver.restore)
    def restore(self, model_path):
        """Stands in for ``tf.train.Saver().restore(sess, model_path)`` (run_nanotensor.py:337-338)."""
        variables = ckpt.load_checkpoint(model_path)
        self.set_weights(ckpt.weights_for_network(variables, self.network, self.model_name))
        return self

    def set_weights(self, weights):
        self._weights = weights
        if self._model is not None:
            self._model.close()
        self._model = Model(self.network, weights, self.graph_name, self.seq_len, self.len_x, self.len_y,
                            device=self.device)
        return self

    @property
    def model(self):
        if self._model is None:
            raise RuntimeError("no weights loaded: call restore(model_path) or set_weights(...) first")
        return self._model

    # ------------------------------------------------------------------ sess.run stand-ins
    def run(self, fetches):
        """Evaluate ``fetches`` on the NEXT batch of the dataset iterator."""
        if self.dataset is None:
            raise RuntimeError("run() needs a dataset; use run_batch(x, seq_len) with explicit arrays")
        batch = self.dataset.iterator.get_next()
        x, seq = batch[0], batch[1]
        y = batch[2] if len(batch) > 2 else None
        return self.run_batch(x, seq, fetches=fetches, y=y)

    def run_batch(self, x, seq_len, fetches=None, y=None):
        """One reference batch (= one BN group)."""
        single = isinstance(fetches, Fetch)
        flist = [fetches] if single else list(fetches if fetches is not None else [self.prediction])
        flat = []
        for f in flist:      # the driver passes [self.inference_opts] (a nested list)
            flat.extend(f if isinstance(f, (list, tuple)) else [f])
        want_logits = any(f.name in ("output_layer", "neg_sum_logits", "beam_prediction") for f in flat)
        x = np.ascontiguousarray(x, dtype=np.float32)
        seq_len = np.ascontiguousarray(seq_len, dtype=np.int32)
        n = len(x)
        goff = np.array([0, n], dtype=np.int32)
        labels, llen, logits = self.model.forward_host(x, seq_len, goff, want_logits=want_logits)
        vals = []
        for f in flat:
            vals.append(self._value(f.name, x, seq_len, labels, llen, logits, y))
        if single:
            return vals[0]
        return vals

    def run_job(self, x, seq_len, group_off, want_logits=False):
        """Whole job (many batches) in BN-group-aligned chunks.  Returns (prediction batches as the
        reference would emit them one sess.run at a time, logits or None)."""
        labels, llen, logits = self.model.forward_host(x, seq_len, group_off, want_logits=want_logits)
        preds = []
        for g in range(len(group_off) - 1):
            a, b = int(group_off[g]), int(group_off[g + 1])
            preds.append(self._prediction(labels, llen, a, b))
        return preds, logits

    def _prediction(self, labels, llen, a, b):
        raise NotImplementedError

    def _value(self, name, x, seq_len, labels, llen, logits, y):
        n = len(x)
        if name == "prediction":
            return self._prediction(labels, llen, 0, n)
        if name == "output_layer":
            return self._shape_logits(logits, n)
        if name == "batch_size":
            return n
        if name == "sequence_length":
            return seq_len
        if name == "x":
            return x.reshape(n, self.seq_len, -1) if self.graph_name == "CtcLoss" else x
        if name == "y":
            return y
        if name == "accuracy":
            return self._accuracy(self._prediction(labels, llen, 0, n), y)
        raise KeyError(name)

    def _shape_logits(self, logits, n):
        return logits

    def _accuracy(self, prediction, y):
        raise NotImplementedError


class CrossEntropy(BuildGraph):
    """network.py:452-473: dense output [B*T, len_y], prediction = tf.argmax(output_layer, 1)."""

    graph_name = "CrossEntropy"

    def __init__(self, x_iterator=None, y_iterator=None, seq_iterator=None, network=None, learning_rate=0, seq_len=0,
                 forget_bias=5.0, reuse=None, y_shape=None, x_shape=None,
                 seq_shape=None, len_x=1, len_y=1, mode=0, dataset=None, summary_name="Training", **kw):
        if dataset:
            len_x, len_y, seq_len, mode = dataset.len_x, dataset.len_y, dataset.seq_len, dataset.mode
        super(CrossEntropy, self).__init__(
            x_iterator=x_iterator, y_iterator=y_iterator, seq_iterator=seq_iterator, network=network,
            learning_rate=learning_rate, seq_len=seq_len, forget_bias=forget_bias, reuse=reuse, mode=mode,
            y_shape=y_shape, x_shape=x_shape, seq_shape=seq_shape, len_x=len_x, len_y=len_y,
            summary_name=summary_name, dataset=dataset, **kw)

    def _prediction(self, labels, llen, a, b):
        T = self.seq_len
        return labels[a * T:b * T].astype(np.int64)

    def _accuracy(self, prediction, y):
        """network.py:213-217: mean(argmax(logits) == argmax(y)); y may be one-hot or class indices."""
        y = np.asarray(y)
        if y.ndim == 3:
            y = y.argmax(axis=2)
        return float(np.mean(prediction == y.reshape(-1)))


class CtcLoss(BuildGraph):
    """network.py:476-533: blank class appended (len_y + 1), x reshaped [B,T,1], output
    [B,T,len_y+1], prediction = dense(ctc_greedy_decoder(merge_repeated=True), default -1)."""

    graph_name = "CtcLoss"

    def __init__(self, dataset=None, x_iterator=None, y_iterator=None, seq_iterator=None, network=None,
                 learning_rate=0, seq_len=0, forget_bias=5.0, reuse=None, training=True,
                 y_shape=None, x_shape=None, seq_shape=None, len_x=1, len_y=1, summary_name="Training",
                 mode=2, **kw):
        if dataset:
            len_x, len_y, seq_len, mode = dataset.len_x, dataset.len_y, dataset.seq_len, dataset.mode
        len_y = len_y + 1                       # network.py:495
        if len_x != 1:
            raise Usage("CtcLoss graphs take one sample per timestep (x is reshaped [-1, seq_len, 1])")
        super(CtcLoss, self).__init__(
            x_iterator=x_iterator, y_iterator=y_iterator, seq_iterator=seq_iterator, network=network,
            learning_rate=learning_rate, seq_len=seq_len, forget_bias=forget_bias, reuse=reuse, mode=mode,
            y_shape=y_shape, x_shape=x_shape, seq_shape=seq_shape, len_x=len_x, len_y=len_y,
            summary_name=summary_name, dataset=dataset, **kw)
        self.neg_sum_logits = Fetch(self, "neg_sum_logits")
        self.beam_prediction = Fetch(self, "beam_prediction")       # (decoded list, log_probabilities)
        self.beam_width, self.top_paths = 100, 1

    def _prediction(self, labels, llen, a, b):
        return ctc_dense(labels[a:b], llen[a:b])

    def _value(self, name, x, seq_len, labels, llen, logits, y):
        """Two more fetches next to the reference's ``prediction``: the decoder's second output, which the reference
        drops (``tf.nn.ctc_greedy_decoder(...)[1]``, network.py:531), and Chiron's beam-search prediction
        (``tf.nn.ctc_beam_search_decoder(merge_repeated=False, top_paths=self.top_paths)``, chiron_rcnn_train.py:66-67)."""
        if name in ("neg_sum_logits", "beam_prediction"):
            from . import ctc
            lg = logits.reshape(len(x), self.seq_len, self.len_y)
            if name == "neg_sum_logits":
                return ctc.ctc_greedy_decoder(lg, seq_len, merge_repeated=True, time_major=False, device=str(self.model.device))[1]
            return ctc.ctc_beam_search_decoder(lg, seq_len, beam_width=self.beam_width, top_paths=self.top_paths,
                                               merge_repeated=False, time_major=False, device=str(self.model.device))
        return super(CtcLoss, self)._value(name, x, seq_len, labels, llen, logits, y)

    def _shape_logits(self, logits, n):
        return logits.reshape(n, self.seq_len, self.len_y)

    def _accuracy(self, prediction, y):
        raise NotImplementedError("edit-distance accuracy of the CTC graphs needs labelled data "
                                  "(outside the forward path)")
