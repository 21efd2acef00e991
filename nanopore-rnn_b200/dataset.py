"""Input pipelines with the reference's names, constructor kwargs and batch semantics
(nanotensor/dataset.py of the reference), without tf.data: windows are produced as numpy arrays
and handed to the engine either batch by batch (``iterator.get_next()``, the reference's
sess.run granularity) or as one BN-group-annotated job (``inference_job``), which is what the
B200 path consumes.

Batch semantics kept (they define the BN groups of the residual layers):
* inference (mode 2): consecutive ``batch_size`` windows of ONE read, last batch short
  (dataset.py:374-377, 435-445); windows every ``step`` samples, zero padded, ``seq_len`` = valid
  samples; the read is z-scored as a whole (chiron/chiron/chiron_input.py:220-230).
* NumpyEventData test mode (mode 1): ``len(data)//seq_len`` full windows per ``.npy`` file, tail
  dropped, consecutive ``batch_size`` windows over the concatenation of all files
  (dataset.py:490-517, 532-534).
"""
import collections
import difflib
import logging as log
import os

import numpy as np

from .engine import group_offsets


class OutOfRangeError(Exception):
    """End of the dataset (stands in for tf.errors.OutOfRangeError, run_nanotensor.py:348)."""


def read_signal(file_path, normalize=True):
    """Whitespace separated integers -> float64 array, z-scored over the whole read
    (population std) when ``normalize`` (chiron/chiron/chiron_input.py:220-230)."""
    with open(file_path, "r") as fh:
        sig = np.array(fh.read().split(), dtype=np.int64)
    if len(sig) == 0:
        raise ValueError("empty signal file %s" % file_path)
    sig = sig.astype(np.float64)
    if normalize:
        sig = (sig - sig.mean()) / sig.std()
    return sig


def read_signal_raw(file_path):
    """The raw int16 samples of a ``.signal`` file (input of the device-side packer)."""
    with open(file_path, "r") as fh:
        sig = np.array(fh.read().split(), dtype=np.int64)
    if len(sig) == 0:
        raise ValueError("empty signal file %s" % file_path)
    if sig.min() < -32768 or sig.max() > 32767:
        raise ValueError("signal file %s does not fit int16" % file_path)
    return sig.astype(np.int16)


def read_signals_gpu(paths, stream=None):
    """The text parse of ``read_signal`` for many ``.signal`` files in one device pass (``ntb_parse_signal_host``; chiron
    /chiron/chiron_input.py:220-226): returns (int16 samples of all files back to back, int64 per-file offsets) -- the
    inputs of ``ntb_pack_windows``, which does the z-score and the windows.  No CPU fallback."""
    import ctypes as C
    from . import _lib
    blobs = []
    for path in paths:
        with open(path, "rb") as fh:
            blobs.append(fh.read())
    file_off = np.zeros(len(blobs) + 1, dtype=np.int64)
    file_off[1:] = np.cumsum([len(b) for b in blobs])
    text = np.frombuffer(b"".join(blobs), dtype=np.uint8)
    capacity = int(file_off[-1]) // 2 + len(blobs) + 1          # a sample needs a digit and a separator
    samples = np.empty(capacity, dtype=np.int16)
    read_off = np.zeros(len(blobs) + 1, dtype=np.int64)
    lib = _lib.load()
    _lib.check(lib.ntb_parse_signal_host(text.ctypes.data_as(C.c_void_p), file_off.ctypes.data_as(C.c_void_p), len(blobs),
                                         int(file_off[-1]), capacity, samples.ctypes.data_as(C.c_void_p),
                                         read_off.ctypes.data_as(C.c_void_p), C.c_void_p(stream or 0)))
    return samples[:read_off[-1]], read_off


class _Iterator(object):
    """Initialisable batch iterator (stands in for dataset.make_initializable_iterator())."""

    def __init__(self, make_batches):
        self._make = make_batches
        self._it = None

    def initializer(self, **feed):
        self._it = iter(self._make(**feed))
        return self

    def get_next(self):
        if self._it is None:
            self.initializer()
        try:
            return next(self._it)
        except StopIteration:
            raise OutOfRangeError()


class CreateDataset(object):
    """Base class: argument checks and attributes as dataset.py:31-107."""

    def __init__(self, mode=0, x_shape=list(), y_shape=list(), sequence_shape=list(), batch_size=10,
                 seq_len=10, len_y=0, len_x=0, n_epochs=5, verbose=False,
                 shuffle_buffer_size=10000, prefetch_buffer_size=100, inference_output_dir="path",
                 file_list="list"):
        # same argument checks (and messages) as dataset.py:50-64
        for name, val, typ in (("mode", mode, int), ("x_shape", x_shape, list), ("y_shape", y_shape, list),
                               ("sequence_shape", sequence_shape, list), ("batch_size", batch_size, int),
                               ("seq_len", seq_len, int), ("len_y", len_y, int), ("len_x", len_x, int),
                               ("verbose", verbose, bool), ("n_epochs", n_epochs, int),
                               ("shuffle_buffer_size", shuffle_buffer_size, int),
                               ("prefetch_buffer_size", prefetch_buffer_size, int)):
            assert type(val) is typ, "{0} is not {1}: type({0}) = {2}".format(
                name if name != "mode" else "mode option", typ.__name__, type(val))
        if mode == 0:
            raise NotImplementedError("training pipelines (mode 0) are outside the B200 forward path")
        self.seq_len = seq_len
        self.batch_size = batch_size
        self.n_epochs = n_epochs
        self.verbose = verbose
        self.len_y = len_y
        self.len_x = len_x
        self.x_shape = x_shape
        self.y_shape = y_shape
        self.sequence_shape = sequence_shape
        self.shuffle_buffer_size = shuffle_buffer_size
        self.prefetch_buffer_size = prefetch_buffer_size
        self.mode = mode
        self.file_list = file_list
        self.file_path = self.file_list[0] if len(self.file_list) else None
        self.inference_output_dir = inference_output_dir
        log.info("Batch Size = {}".format(self.batch_size))
        log.info("Sequence Length = {}".format(self.seq_len))
        log.info("Size of input vector = {}".format(self.len_x))
        log.info("Size of label vector = {}".format(self.len_y))
        self.training_labels = collections.namedtuple('training_data', ['input', 'seq_len', 'label'])
        self.inference_labels = collections.namedtuple('inference_data', ['input', 'seq_len'])
        self.data = self.load_data() if self.mode == 1 else None
        self.iterator = self.create_iterator()

    # -- tf.data stand-ins ------------------------------------------------------------------
    def create_iterator(self):
        return _Iterator(self._batches)

    def _batches(self, **feed):
        bs = self.batch_size
        if self.mode == 2:
            x, lens = self.load_file_windows(self.file_path)
            for s in range(0, len(x), bs):
                yield x[s:s + bs], lens[s:s + bs]
        else:
            data = self.data
            n = len(data.input)
            for s in range(0, n, bs):
                yield (np.asarray(data.input[s:s + bs], dtype=np.float32),
                       np.asarray(data.seq_len[s:s + bs], dtype=np.int32), data.label[s:s + bs])

    @staticmethod
    def test_numpy_files(path_list):
        """dataset.py:141-155: split into loadable / unloadable ``.npy`` files."""
        assert type(path_list) is list, "path_list is not list: type(path_list) = {}".format(type(path_list))
        good, bad = [], []
        for f in path_list:
            try:
                np.load(f, allow_pickle=True, encoding="latin1")
                good.append(f)
            except Exception:   # unreadable / not an npy file (the reference catches IOError)
                bad.append(f)
        return good, bad

    @staticmethod
    def padding(x, L, padding_list=None, pad=0):
        """dataset.py:157-169."""
        len_x = len(x)
        assert len_x <= L, "Length of vector x is larger than the padding length"
        assert type(pad) is int, "Pad should be an integer"
        zero_n = L - len_x
        x = list(x)
        if padding_list is None:
            return x + [pad] * zero_n
        elif len(padding_list) < zero_n:
            return x + list(padding_list) + [pad] * (zero_n - len(padding_list))
        return x + list(padding_list[0:zero_n])

    # -- to be provided by subclasses -----------------------------------------------------------
    def load_data(self):
        raise NotImplementedError

    def load_file_windows(self, path):
        raise NotImplementedError

    def process_output(self, graph_output, input_path):
        raise NotImplementedError

    def load_data_inference(self):
        """Generator of (padded window, valid length) for ``self.file_path`` (dataset.py:435-445)."""
        x, lens = self.load_file_windows(self.file_path)
        for i in range(len(x)):
            yield self.inference_labels(input=x[i], seq_len=lens[i])

    # -- the B200 job form ------------------------------------------------------------------
    def inference_job(self, files=None):
        """All windows of ``files`` as one job: x, lens, BN group offsets (= the reference's
        batches, never spanning two reads) and per-read window offsets."""
        files = list(files) if files is not None else list(self.file_list)
        xs, ls, counts = [], [], []
        for f in files:
            x, lens = self.load_file_windows(f)
            xs.append(x)
            ls.append(lens)
            counts.append(len(x))
        shape = (0, self.seq_len) + ((self.len_x,) if self.len_x > 1 else ())
        x = np.concatenate(xs) if xs else np.zeros(shape, np.float32)
        lens = np.concatenate(ls) if ls else np.zeros((0,), np.int32)
        read_off = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
        return x, lens, group_offsets(counts, self.batch_size), read_off


def windows_from_signal(sig, seq_len, step, start_index):
    """dataset.py:435-445 on an already normalised read: float32 [n, seq_len], int32 [n]."""
    sig = np.asarray(sig, dtype=np.float64)[start_index:]
    n = len(sig)
    starts = np.arange(0, n, step)
    x = np.zeros((len(starts), seq_len), dtype=np.float32)
    lens = np.minimum(seq_len, n - starts).astype(np.int32)
    for i, s in enumerate(starts):
        x[i, :lens[i]] = sig[s:s + lens[i]]
    return x, lens


class FullSignalSequence(CreateDataset):
    """Raw-signal windows for the CtcLoss graphs (dataset.py:305-445)."""

    def __init__(self, file_list, mode=0, batch_size=10, verbose=True, seq_len=100,
                 n_epochs=5, shuffle_buffer_size=10000, prefetch_buffer_size=100, step=300, start_index=0,
                 inference_output_dir="path", alphabet=5, max_event_len=50):
        self.file_list = file_list
        self.len_y = alphabet
        self.len_x = 1
        self.kmer = 1
        self.step = step
        self.start_index = start_index
        self.max_event_len = max_event_len
        super(FullSignalSequence, self).__init__(
            mode=mode, x_shape=[None, seq_len], y_shape=[None, None], sequence_shape=[None],
            batch_size=batch_size, seq_len=seq_len, len_y=self.len_y, len_x=self.len_x, n_epochs=n_epochs,
            verbose=verbose, shuffle_buffer_size=shuffle_buffer_size,
            prefetch_buffer_size=prefetch_buffer_size, inference_output_dir=inference_output_dir,
            file_list=file_list)

    def load_data(self):
        raise NotImplementedError("labelled .signal/.label loading (training/test of the CTC graphs) "
                                  "is outside the forward path")

    def load_file_windows(self, path):
        return windows_from_signal(read_signal(path, normalize=True), self.seq_len, self.step, self.start_index)

    def process_output(self, graph_output, input_path):
        """dataset.py:421-432: decoded label rows -> bases -> consensus -> ``<name>.fasta``."""
        name = os.path.splitext(os.path.basename(input_path))[0]
        fasta_out_path = os.path.join(self.inference_output_dir, name + ".fasta")
        consensus = assemble_read(graph_output)
        with open(fasta_out_path, "w+") as fasta_f:
            fasta_f.write(">{}\n{}\n".format(name, consensus))
        return fasta_out_path


class MotifSequence(FullSignalSequence):
    """Same inference pipeline as FullSignalSequence (dataset.py:188-302: identical
    load_data_inference / process_output); ``blank`` selects the 3-symbol motif alphabet."""

    def __init__(self, file_list, mode=0, batch_size=10, verbose=True, seq_len=100,
                 n_epochs=5, shuffle_buffer_size=10000, prefetch_buffer_size=100, blank=True,
                 inference_output_dir="path", step=300, start_index=0):
        self.blank = blank
        super(MotifSequence, self).__init__(
            file_list, mode=mode, batch_size=batch_size, verbose=verbose, seq_len=seq_len, n_epochs=n_epochs,
            shuffle_buffer_size=shuffle_buffer_size, prefetch_buffer_size=prefetch_buffer_size, step=step,
            start_index=start_index, inference_output_dir=inference_output_dir, alphabet=3 if blank else 5)


class NumpyEventData(CreateDataset):
    """Event-feature windows from ``.npy`` object arrays for the CrossEntropy graph
    (dataset.py:448-543)."""

    def __init__(self, file_list, mode=0, batch_size=10, verbose=True, seq_len=100,
                 n_epochs=5, shuffle_buffer_size=10000, prefetch_buffer_size=100, inference_output_dir="path"):
        self.file_list, self.bad_files = self.test_numpy_files(list(file_list))
        self.num_files = len(self.file_list)
        assert self.num_files >= 1, "There are no passing npy files to read into queue"
        if self.bad_files:
            log.warning("{} files were unable to be loaded using np.load()".format(len(self.bad_files)))
        data = np.load(self.file_list[0], allow_pickle=True, encoding="latin1")
        self.len_x = len(data[0][0])
        self.len_y = len(data[0][1])
        super(NumpyEventData, self).__init__(
            mode=mode, x_shape=[None, seq_len, self.len_x], y_shape=[None, seq_len, self.len_y],
            sequence_shape=[None], batch_size=batch_size, seq_len=seq_len, len_y=self.len_y, len_x=self.len_x,
            n_epochs=n_epochs, verbose=verbose, shuffle_buffer_size=shuffle_buffer_size,
            prefetch_buffer_size=prefetch_buffer_size, inference_output_dir=inference_output_dir,
            file_list=self.file_list)

    def _file_windows(self, path, with_labels):
        data = np.load(path, allow_pickle=True, encoding="latin1")
        nb = len(data) // self.seq_len
        feats = np.zeros((nb, self.seq_len, self.len_x), dtype=np.float32)
        labels = np.zeros((nb, self.seq_len), dtype=np.int64) if with_labels else None
        for b in range(nb):
            chunk = data[b * self.seq_len:(b + 1) * self.seq_len]
            feats[b] = np.vstack(chunk[:, 0])
            if with_labels:
                labels[b] = np.argmax(np.vstack(chunk[:, 1]), axis=1)
        return feats, labels

    def load_data(self):
        """dataset.py:490-517.  Labels are kept as class indices (argmax of the one-hot rows), which
        is all the accuracy function uses (network.py:213-217)."""
        xs, ys = [], []
        for f in self.file_list:
            x, y = self._file_windows(f, True)
            xs.append(x)
            ys.append(y)
        x = np.concatenate(xs)
        return self.training_labels(input=x, seq_len=np.full((len(x),), self.seq_len, dtype=np.int32),
                                    label=np.concatenate(ys))

    def load_file_windows(self, path):
        x, _ = self._file_windows(path, False)
        return x, np.full((len(x),), self.seq_len, dtype=np.int32)

    def process_output(self, graph_output, input_path):
        for x in graph_output:
            print(x)

    def test_job(self):
        """All mode-1 windows as one job; groups = consecutive batch_size windows over the
        concatenation (dataset.py:532-534)."""
        n = len(self.data.input)
        return (np.asarray(self.data.input, np.float32), np.asarray(self.data.seq_len, np.int32),
                group_offsets([n], self.batch_size), np.asarray(self.data.label))


# ---------------------------------------------------------------------------------------------
# post-processing: label rows -> bases -> consensus read
# ---------------------------------------------------------------------------------------------
_BASES = ("A", "C", "G", "T", "E")
_BASE_INDEX = {b: i for i, b in enumerate(_BASES)}


def index2base(read, blank=False):
    """nanotensor/trim_signal.py:116-122 (SignalLabel.index2base): class index -> letter, -1 -> ''."""
    table = ("", "C", "E") if blank else _BASES
    return "".join(table[int(k)] for k in read if int(k) != -1)


def simple_assembly(bpreads):
    """Overlap-place consecutive window calls and count votes per position; restates the
    behaviour of chiron/chiron/utils/easy_assembler.py:11-37: each read is shifted against its
    predecessor by the offset of their longest common block (difflib), negative starts are
    clipped, and a [5, L] vote matrix is returned."""
    votes = collections.defaultdict(lambda: np.zeros(5))
    cursor, length = 0, 0
    prev = None
    for read in bpreads:
        if prev is not None:
            blocks = difflib.SequenceMatcher(None, prev, read).get_matching_blocks()
            best = max(blocks, key=lambda m: m[2])
            cursor += best[0] - best[1]
        start, seg = cursor, read
        if start < 0:
            seg, start = seg[-start:], 0
        for i, ch in enumerate(seg):
            votes[start + i][_BASE_INDEX[ch.upper()]] += 1
        if prev is not None:   # the reference only extends the length from the 2nd read on
            length = max(length, cursor + len(read))
        prev = read
    out = np.zeros((5, max(length, 0)))
    for pos, v in votes.items():
        if pos < out.shape[1]:
            out[:, pos] = v
    return out


def assemble_read(label_batches, blank=False):
    """dataset.py:421-432 without the file write: batches of label rows -> consensus string."""
    reads = []
    for batch in label_batches:
        reads.extend(index2base(r, blank=blank) for r in batch)
    if not reads:
        return ""
    cons = simple_assembly(reads)
    if cons.shape[1] == 0:
        return ""
    return index2base(np.argmax(cons, axis=0))
