"""Host-side engine: owns the GPU-resident repacked weights (through the C ABI), plans
BN-group-aligned chunks of windows and drives ``ntb_model_forward`` / ``_forward_host``.

torch is used for device memory, streams and pinned host buffers only."""
import ctypes as C
import os
import weakref

import numpy as np
import torch

from . import _lib
from ._lib import (NTB_GRAPH_CROSS_ENTROPY, NTB_GRAPH_CTC, NTB_LAYER_BLSTM, NTB_LAYER_CHIRON_FNN,
                   NTB_LAYER_RESIDUAL, NtbLayer, NtbModelDesc)

GRAPHS = {"CrossEntropy": NTB_GRAPH_CROSS_ENTROPY, "CtcLoss": NTB_GRAPH_CTC}
# recurrence tile = 128 windows x 2 directions; 74 tiles x 2 = one wave of 148 SMs
_WAVE_WINDOWS = 128 * 74


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def group_offsets(read_window_counts, batch_size):
    """BN groups = the reference's batches: consecutive ``batch_size`` windows of ONE read, last
    batch of a read short (dataset.py:374-377 ``dataset.batch(batch_size)`` per file,
    run_nanotensor.py:339-347).  Returns int32 offsets [n_groups+1]."""
    offs = [0]
    pos = 0
    for n in read_window_counts:
        n = int(n)
        full, rem = divmod(n, batch_size)
        sizes = [batch_size] * full + ([rem] if rem else [])
        for s in sizes:
            pos += s
            offs.append(pos)
    return np.asarray(offs, dtype=np.int32)


def plan_chunks(group_off, max_windows, align=1):
    """Split [0, n_groups) into runs of whole groups with <= max_windows windows each, preferring
    a multiple of one recurrence wave (128 windows x 74 tiles x 2 directions = 148 CTAs).

    ``align`` (Model.chunk_align()): cuts are placed where the window count so far is a multiple of it whenever
    such a group boundary exists inside the limit.  The conv epilogue sums the batch-norm statistics in 32-row
    blocks counted from the start of a chunk; with aligned cuts those blocks -- hence every bit of the result --
    are the same however the job is cut (unaligned cuts differ by fp32 rounding of the block sums, ~1e-7)."""
    group_off = np.asarray(group_off, dtype=np.int64)
    n_groups = len(group_off) - 1
    total = int(group_off[-1])
    if total <= max_windows:
        return [(0, n_groups)]
    target = max_windows
    if max_windows >= _WAVE_WINDOWS:
        target = (max_windows // _WAVE_WINDOWS) * _WAVE_WINDOWS
    chunks = []
    g = 0
    while g < n_groups:
        limit = group_off[g] + target
        h = int(np.searchsorted(group_off, limit, side="right")) - 1
        if align > 1 and h > g and h < n_groups:
            ok = np.nonzero(group_off[g + 1:h + 1] % align == 0)[0]
            if len(ok):
                h = g + 1 + int(ok[-1])
        if h <= g:
            h = g + 1
            if group_off[h] - group_off[g] > max_windows:
                raise ValueError("a single BN group (%d windows) exceeds the chunk size %d"
                                 % (group_off[h] - group_off[g], max_windows))
        chunks.append((g, h))
        g = h
    return chunks


class Model:
    """GPU-resident NanoTensor network (config ``network`` list + TF-layout weights)."""

    def __init__(self, network, weights, graph, seq_len, len_x, n_classes, device="cuda:0"):
        self.lib = _lib.load()
        if graph not in GRAPHS:
            raise ValueError("graph must be one of %s" % sorted(GRAPHS))
        self.network = [dict(l) for l in network]
        self.graph = graph
        self.T, self.len_x, self.C = int(seq_len), int(len_x), int(n_classes)
        self.device = torch.device(device)
        if self.device.type != "cuda" or not torch.cuda.is_available():
            raise _lib.NtbError(-2, "nanotensor_b200 needs a CUDA sm_100 device; there is no CPU fallback")
        self.has_residual = any(l["type"] == "residual_layer" for l in network)
        keep = []   # host arrays referenced by the desc must outlive ntb_model_create

        def arr(a):
            a = np.ascontiguousarray(a, dtype=np.float32)
            keep.append(a)
            return _fp(a)

        layers = (NtbLayer * len(network))()
        dim = self.len_x
        for i, (l, w) in enumerate(zip(network, weights["layers"])):
            L = layers[i]
            L.in_dim = dim
            if l["type"] == "residual_layer":
                L.type, L.size, L.batchnorm = NTB_LAYER_RESIDUAL, int(l["out_channel"]), int(bool(l.get("batchnorm")))
                order = ["conv1", "conv1_bn_scale", "conv1_bn_offset", "conv2a", "conv2a_bn_scale",
                         "conv2a_bn_offset", "conv2b", "conv2b_bn_scale", "conv2b_bn_offset", "conv2c",
                         "conv2c_bn_scale", "conv2c_bn_offset"]
                for k, name in enumerate(order):
                    if name in w:
                        L.w[k] = arr(w[name])
                dim = L.size
            elif l["type"] == "blstm":
                if not l.get("concat", True):
                    raise ValueError("blstm layer %r: concat must be true (network.py:284-287)" % l.get("name"))
                L.type, L.size, L.forget_bias = NTB_LAYER_BLSTM, int(l["size"]), float(l["bias"])
                for k, name in enumerate(["kernel_fw", "bias_fw", "kernel_bw", "bias_bw"]):
                    L.w[k] = arr(w[name])
                if w["kernel_fw"].shape != (dim + L.size, 4 * L.size):
                    raise ValueError("blstm layer %r: kernel shape %s, expected %s"
                                     % (l.get("name"), w["kernel_fw"].shape, (dim + L.size, 4 * L.size)))
                dim = 2 * L.size
            elif l["type"] == "chiron_fnn":
                L.type, L.size = NTB_LAYER_CHIRON_FNN, dim // 2
                L.w[0], L.w[1] = arr(w["weights"]), arr(w["bias"])
                dim = dim // 2
            else:
                raise ValueError("layer type %r is not on the supported forward path" % l["type"])
        if tuple(weights["out_w"].shape) != (dim, self.C):
            raise ValueError("output layer weights %s, expected %s" % (weights["out_w"].shape, (dim, self.C)))
        self.final_dim = dim
        desc = NtbModelDesc(GRAPHS[graph], self.T, self.len_x, self.C, len(network), layers,
                            arr(weights["out_w"]), arr(weights["out_b"]))
        nbytes = self.lib.ntb_model_weight_bytes(C.byref(desc))
        if nbytes < 0:
            _lib.check(int(nbytes))
        with torch.cuda.device(self.device):
            self._weights = torch.empty(int(nbytes), dtype=torch.uint8, device=self.device)
            handle = C.c_void_p()
            _lib.check(self.lib.ntb_model_create(C.byref(desc), _ptr(self._weights), nbytes,
                                                 _stream_ptr(self.device), C.byref(handle)))
        self._h = handle
        self._ws = None
        self._chunk_cap = None    # max_chunk_windows() with the default budget, computed once
        self._stage = None        # (key, two sets of device staging buffers) of the overlapped host path
        self._streams = None
        del keep

    def close(self):
        if getattr(self, "_h", None):
            self.lib.ntb_model_destroy(self._h)
            self._h = None
            self._ws = self._stage = None

    __del__ = close

    # ------------------------------------------------------------------ workspace
    def workspace_bytes(self, n_windows, n_groups, host=False, want_logits=False):
        if host:
            return int(self.lib.ntb_model_workspace_bytes_host(self._h, n_windows, n_groups, int(want_logits)))
        return int(self.lib.ntb_model_workspace_bytes(self._h, n_windows, n_groups))

    def _workspace(self, nbytes):
        if self._ws is None or self._ws.numel() < nbytes:
            self._ws = None
            self._ws = torch.empty(int(nbytes), dtype=torch.uint8, device=self.device)
        return self._ws

    def chunk_align(self):
        """Window-count multiple at which chunk cuts keep the fused BN statistics bit-identical (plan_chunks)."""
        if not any(l["type"] == "residual_layer" for l in self.network) or self.T < 32:
            return 1
        a = 1
        for tp in (self.T, self.T + 2):          # plane rows per window: 1x1 convs, padded 1x3 conv
            a = int(np.lcm(a, 32 // int(np.gcd(tp, 32))))
        return a

    def max_chunk_windows(self, budget_bytes=None, groups_per_window=0.1):
        """Largest chunk whose workspace fits ``budget_bytes`` (default: 60 % of free HBM)."""
        default = budget_bytes is None
        if default:
            # cudaMemGetInfo costs ~9 ms: ask once while no workspace is cached (a cached one already fits its chunk)
            if self._chunk_cap is not None:
                return self._chunk_cap
            free, _ = torch.cuda.mem_get_info(self.device)
            budget_bytes = int(free * 0.6)
        per = self.workspace_bytes(4096, int(4096 * groups_per_window) + 1, host=True) / 4096.0
        n = int(budget_bytes / per)
        n = max(64, min(n, 16 * _WAVE_WINDOWS))
        if default:
            self._chunk_cap = n
        return n

    # ------------------------------------------------------------------ forward, device-resident inputs
    def forward_device(self, x, lens, group_off=None, want_logits=False, labels_out=None, label_len_out=None):
        """One chunk with inputs already in HBM.  x [n,T(,len_x)] f32, lens [n] i32, group_off
        [n_groups+1] i32 (all cuda tensors).  Returns (labels i32, label_len i32 | None, logits | None)."""
        n = int(x.shape[0])
        ng = int(group_off.numel()) - 1 if group_off is not None else 0
        if self.has_residual and group_off is None:
            raise ValueError("networks with residual layers need group_off (BN groups)")
        rows = n * self.T
        with torch.cuda.device(self.device):
            ws = self._workspace(self.workspace_bytes(n, ng))
            labels = labels_out if labels_out is not None else torch.empty(max(rows, 1), dtype=torch.int32, device=self.device)
            llen = None
            if self.graph == "CtcLoss":
                llen = label_len_out if label_len_out is not None else torch.empty(max(n, 1), dtype=torch.int32, device=self.device)
            logits = torch.empty((rows, self.C), dtype=torch.float32, device=self.device) if want_logits else None
            _lib.check(self.lib.ntb_model_forward(
                self._h, _ptr(x), _ptr(lens), _ptr(group_off), n, ng, _ptr(ws), ws.numel(),
                _ptr(logits), _ptr(labels), _ptr(llen), _stream_ptr(self.device)))
        return labels[:rows], (llen[:n] if llen is not None else None), logits

    # ------------------------------------------------------------------ forward, host buffers (end to end)
    def forward_host(self, x, lens, group_off=None, want_logits=False, max_windows=None):
        """Whole job from HOST arrays (numpy or CPU torch tensors; pinned is faster): chunked on BN
        group boundaries, host->device copies and device->host label reads inside.  Returns
        (labels int32 [n*T] (CrossEntropy) or [n,T] (CtcLoss, -1 padded), label_len int32 [n] | None,
        logits float32 [n*T, C] | None) as numpy arrays."""
        x = _as_np(x, np.float32)
        lens = _as_np(lens, np.int32)
        n = int(x.shape[0])
        if group_off is None:
            if self.has_residual:
                raise ValueError("networks with residual layers need group_off (BN groups)")
            step = max_windows or self.max_chunk_windows()
            group_off = np.arange(0, n + step, step, dtype=np.int64).clip(max=n)
            group_off = np.unique(group_off).astype(np.int32)
            if n == 0:
                group_off = np.zeros(1, dtype=np.int32)
            pass_groups = False
        else:
            group_off = _as_np(group_off, np.int32)
            pass_groups = True
            if len(group_off) < 1 or group_off[0] != 0 or group_off[-1] != n or np.any(np.diff(group_off) < 0):
                raise ValueError("group_off must ascend from 0 to the number of windows (%d)" % n)
        if len(lens) != n or (n and (lens.min() < 0 or lens.max() > self.T)):
            raise ValueError("lens must hold one value in [0, seq_len] per window")
        T, Cc = self.T, self.C
        # results land in pinned host memory (torch's caching host allocator recycles the blocks)
        labels = _pinned((n * T,), torch.int32)
        llen = _pinned((n,), torch.int32) if self.graph == "CtcLoss" else None
        logits = _pinned((n * T, Cc), torch.float32) if want_logits else None
        if n == 0:
            return self._shape_labels(labels, n), llen, logits
        chunks = plan_chunks(group_off, max_windows or self.max_chunk_windows(), self.chunk_align())
        if len(chunks) >= 2 and os.environ.get("NTB_HOST_PIPELINE", "1") != "0":
            self._forward_host_overlapped(x, lens, group_off, pass_groups, chunks, labels, llen, logits)
            return self._shape_labels(labels, n), llen, logits
        stream = _stream_ptr(self.device)
        with torch.cuda.device(self.device):
            wmax = max(int(group_off[h] - group_off[g]) for g, h in chunks)
            gmax = max(h - g for g, h in chunks)
            ws = self._workspace(self.workspace_bytes(wmax, gmax, host=True, want_logits=want_logits))
            for g, h in chunks:
                w0, w1 = int(group_off[g]), int(group_off[h])
                nw = w1 - w0
                goff = np.ascontiguousarray(group_off[g:h + 1] - group_off[g], dtype=np.int32)
                xs = x[w0:w1]
                ls = lens[w0:w1]
                _lib.check(self.lib.ntb_model_forward_host(
                    self._h, C.c_void_p(xs.ctypes.data), C.c_void_p(ls.ctypes.data),
                    C.c_void_p(goff.ctypes.data) if pass_groups else C.c_void_p(0), nw,
                    (h - g) if pass_groups else 0, _ptr(ws), ws.numel(),
                    C.c_void_p(logits[w0 * T:].ctypes.data) if want_logits else C.c_void_p(0),
                    C.c_void_p(labels[w0 * T:].ctypes.data),
                    C.c_void_p(llen[w0:].ctypes.data) if llen is not None else C.c_void_p(0), stream))
        return self._shape_labels(labels, n), llen, logits

    def _forward_host_overlapped(self, x, lens, group_off, pass_groups, chunks, labels, llen, logits):
        """Several chunks: the kernels of all chunks run back to back on the caller's stream through
        ``ntb_model_forward`` with ONE workspace, while a second stream carries the host->device copy of chunk i+1 and
        the device->host read of chunk i-1 through two sets of staging buffers.  Results are bit-identical to the
        synchronous per-chunk ``ntb_model_forward_host`` calls (same chunks, same kernels)."""
        T, Cc, dev = self.T, self.C, self.device
        want_logits = logits is not None
        wmax = max(int(group_off[h] - group_off[g]) for g, h in chunks)
        gmax = max(h - g for g, h in chunks)
        xt, lt = torch.from_numpy(_pin(x)), torch.from_numpy(_pin(lens))
        lab_t = torch.from_numpy(labels)
        llen_t = torch.from_numpy(llen) if llen is not None else None
        log_t = torch.from_numpy(logits) if want_logits else None
        with torch.cuda.device(dev):
            ws = self._workspace(self.workspace_bytes(wmax, gmax))
            key = (wmax, gmax, want_logits, tuple(x.shape[1:]))
            if self._stage is None or self._stage[0] != key:
                def bufs():
                    return {"x": torch.empty((wmax,) + tuple(x.shape[1:]), dtype=torch.float32, device=dev),
                            "len": torch.empty(wmax, dtype=torch.int32, device=dev),
                            "goff": torch.empty(gmax + 1, dtype=torch.int32, device=dev),
                            "lab": torch.empty(wmax * T, dtype=torch.int32, device=dev),
                            "llen": torch.empty(wmax, dtype=torch.int32, device=dev),
                            "log": torch.empty((wmax * T, Cc), dtype=torch.float32, device=dev) if want_logits else None}
                self._stage = (key, [bufs(), bufs()])
            stage = self._stage[1]
            if self._streams is None:
                self._streams = [torch.cuda.Stream(dev)]
            copy = self._streams[0]
            main = torch.cuda.current_stream(dev)
            copy.wait_stream(main)
            goffs = [_pin(np.ascontiguousarray(group_off[g:h + 1] - group_off[g], dtype=np.int32)) for g, h in chunks]
            in_ready = [torch.cuda.Event() for _ in chunks]
            out_ready = [torch.cuda.Event() for _ in chunks]
            drained = [torch.cuda.Event() for _ in chunks]        # staging set free again (its D2H has finished)

            def upload(i):
                g, h = chunks[i]
                w0, w1 = int(group_off[g]), int(group_off[h])
                b = stage[i & 1]
                with torch.cuda.stream(copy):
                    if i >= 2:
                        copy.wait_event(out_ready[i - 2])          # chunk i-2 no longer reads / writes this set
                    b["x"][:w1 - w0].copy_(xt[w0:w1], non_blocking=True)
                    b["len"][:w1 - w0].copy_(lt[w0:w1], non_blocking=True)
                    if pass_groups:
                        b["goff"][:h - g + 1].copy_(torch.from_numpy(goffs[i]), non_blocking=True)
                    in_ready[i].record(copy)

            upload(0)
            for i, (g, h) in enumerate(chunks):
                w0, w1 = int(group_off[g]), int(group_off[h])
                nw = w1 - w0
                b = stage[i & 1]
                if i + 1 < len(chunks):
                    upload(i + 1)
                main.wait_event(in_ready[i])
                if i >= 2:
                    main.wait_event(drained[i - 2])                # the labels of chunk i-2 have left this set
                _lib.check(self.lib.ntb_model_forward(
                    self._h, _ptr(b["x"]), _ptr(b["len"]), _ptr(b["goff"]) if pass_groups else C.c_void_p(0), nw,
                    (h - g) if pass_groups else 0, _ptr(ws), ws.numel(), _ptr(b["log"]), _ptr(b["lab"]), _ptr(b["llen"]),
                    C.c_void_p(main.cuda_stream)))
                out_ready[i].record(main)
                with torch.cuda.stream(copy):
                    copy.wait_event(out_ready[i])
                    lab_t[w0 * T:w1 * T].copy_(b["lab"][:nw * T], non_blocking=True)
                    if llen_t is not None:
                        llen_t[w0:w1].copy_(b["llen"][:nw], non_blocking=True)
                    if want_logits:
                        log_t[w0 * T:w1 * T].copy_(b["log"][:nw * T], non_blocking=True)
                    drained[i].record(copy)
            copy.synchronize()
            main.synchronize()

    def _shape_labels(self, labels, n):
        return labels.reshape(n, self.T) if self.graph == "CtcLoss" else labels

    # ------------------------------------------------------------------ introspection
    def last_launches(self):
        return int(self.lib.ntb_model_last_launches(self._h))

    def set_profile(self, on):
        self.lib.ntb_model_set_profile(self._h, int(bool(on)))

    def stage_ms(self):
        buf = (C.c_float * 5)()
        k = self.lib.ntb_model_stage_ms(self._h, buf, 5)
        return dict(zip(["conv_bn", "projection_split", "recurrence", "readout", "projection_gemm"][:k], list(buf)[:k]))


def _pin(a):
    """numpy array -> the same data in page-locked memory (no copy if torch says it already is pinned)."""
    t = torch.from_numpy(a)
    if t.numel() == 0 or t.is_pinned():
        return a
    return t.pin_memory().numpy()


_PIN_POOL = []      # [pinned torch tensor, weakref to the numpy array handed out]: page-locking 100 MB costs ~9 ms


def _pinned(shape, dtype):
    """Page-locked numpy array for results.  Blocks are recycled once the caller has dropped the array (and every
    view of it) that was handed out before: a weak reference to that array tells."""
    count = int(np.prod(shape))
    if count == 0:
        return np.empty(shape, dtype=torch.empty(0, dtype=dtype).numpy().dtype)
    for entry in _PIN_POOL:
        t, ref = entry
        if ref() is None and t.dtype == dtype and t.numel() == count:
            a = t.numpy().reshape(shape)
            entry[1] = weakref.ref(a.base if a.base is not None else a)
            return a
    if len(_PIN_POOL) >= 8:                       # drop the idle blocks before growing without bound
        _PIN_POOL[:] = [e for e in _PIN_POOL if e[1]() is not None]
    t = torch.empty(count, dtype=dtype, pin_memory=True)
    base = t.numpy()
    _PIN_POOL.append([t, weakref.ref(base)])
    return base.reshape(shape)


def _as_np(a, dtype):
    if isinstance(a, torch.Tensor):
        a = a.detach().cpu().numpy()
    return np.ascontiguousarray(a, dtype=dtype)


def ctc_dense(labels, label_len):
    """[n, T] -1-padded device/host labels -> the reference's dense int64 [n, max_decoded]
    (tf.sparse_tensor_to_dense(default_value=-1), network.py:532)."""
    labels = np.asarray(labels)
    width = int(label_len.max()) if len(label_len) else 0
    return labels[:, :width].astype(np.int64)
