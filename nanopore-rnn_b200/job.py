"""The device-resident inference job (SURVEY.md 8(a) rows a9/a11 + 8(e) + 8(f) ranks 1-2): what one tower of
``RunTensorflow.run_model`` does for its share of the reads (run_nanotensor.py:188-202, 339-351), with every
intermediate kept in HBM:

    raw int16 samples (or ``.signal`` text)  --H2D, 2 B/sample-->
    ntb_parse_signal (text only) -> ntb_pack_windows (z-score over the whole read, windows, zero pad:
    dataset.py:435-445, chiron_input.py:220-230) -> ntb_model_forward per BN-group-aligned chunk ->
    ntb_assemble_reads (process_output: index2base + simple_assembly + argmax, dataset.py:421-432)
    --D2H, one byte per called base--> consensus of every read.

Reads are packed in descending window count ("length-sorted packing"), results come back in the caller's order.
Several ranks: ``shard.assign_reads`` gives every rank whole reads (LPT on window counts); ``gather_consensus`` brings
the called reads to rank 0 over NCCL (uint8 buffers in HBM, one device->host copy on rank 0) and restores the
original read order.  No collective sits on the compute path.  torch is the allocator / stream / NCCL plumbing only;
there is no CPU fallback.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from .engine import _pin, _ptr, _stream_ptr, plan_chunks


def window_counts(read_lengths, step, start_index=0):
    """Windows per read: one every ``step`` samples of the read past ``start_index`` (dataset.py:439-441)."""
    n = np.maximum(np.asarray(read_lengths, dtype=np.int64) - int(start_index), 0)
    return -(-n // int(step))


def group_offsets_vec(counts, batch_size):
    """engine.group_offsets without the Python loop: BN groups = consecutive ``batch_size`` windows of one read."""
    counts = np.asarray(counts, dtype=np.int64)
    ng = -(-counts // batch_size)
    total_g = int(ng.sum())
    win_off = np.concatenate([[0], np.cumsum(counts)])
    g_first = np.concatenate([[0], np.cumsum(ng)])[:-1]
    read_of_group = np.repeat(np.arange(len(counts)), ng)
    k = np.arange(total_g) - g_first[read_of_group]
    starts = win_off[:-1][read_of_group] + k * batch_size
    return np.concatenate([starts, [win_off[-1]]]).astype(np.int32)


class JobPlan(object):
    """Host-side table of one rank's share: packing order, per-read sample / window offsets, BN groups, chunks."""

    def __init__(self, read_lengths, seq_len, step, start_index, batch_size, chunk_cap, chunk_align):
        self.lengths = np.asarray(read_lengths, dtype=np.int64)
        self.seq_len, self.step, self.start_index, self.batch = int(seq_len), int(step), int(start_index), int(batch_size)
        counts = window_counts(self.lengths, step, start_index)
        self.order = np.argsort(-counts, kind="stable")              # packed position -> caller's read index
        self.counts = counts[self.order]
        self.read_off = np.concatenate([[0], np.cumsum(self.lengths[self.order])]).astype(np.int64)
        self.win_off = np.concatenate([[0], np.cumsum(self.counts)]).astype(np.int64)
        self.n_reads, self.n_windows = len(self.lengths), int(self.win_off[-1])
        self.total_samples = int(self.read_off[-1])
        self.group_off = group_offsets_vec(self.counts, self.batch) if batch_size > 0 else None
        if self.group_off is not None and self.n_windows:
            self.chunks = plan_chunks(self.group_off, chunk_cap, chunk_align)
        else:
            self.chunks = []
        # valid samples inside windows (the bench's "samples basecalled"): sum of the window lengths
        self.valid_samples = sum_valid(np.maximum(self.lengths - self.start_index, 0), self.step, self.seq_len)


def sum_valid(n, step, seq_len):
    """Sum over windows of min(seq_len, remaining) for reads of ``n`` samples (vectorised)."""
    n = np.asarray(n, dtype=np.int64)
    cnt = -(-n // step)
    last_start = (cnt - 1) * step
    # every window but those near the end is full; windows starting at s have min(seq_len, n - s) samples
    if seq_len <= step:
        full = np.maximum(cnt - 1, 0) * seq_len
        return int((full + np.where(cnt > 0, np.minimum(seq_len, n - last_start), 0)).sum())
    total = 0
    for ni, ci in zip(n.tolist(), cnt.tolist()):                       # overlapping windows: rare path, plain loop
        s = np.arange(ci) * step
        total += int(np.minimum(seq_len, ni - s).sum())
    return total


class ReadJob(object):
    """Runs whole reads through ``model`` (an ``engine.Model`` with the CtcLoss readout) on its device."""

    def __init__(self, model, step=300, start_index=0, batch_size=10, chunk_windows=None):
        if model.graph != "CtcLoss" or model.len_x != 1:
            raise ValueError("ReadJob drives the raw-signal CtcLoss graphs (FullSignalSequence / MotifSequence)")
        self.m, self.lib = model, _lib.load()
        self.T, self.step, self.start_index, self.batch = model.T, int(step), int(start_index), int(batch_size)
        self.chunk_windows = chunk_windows
        self._bufs = {}
        self.last = {}

    # ------------------------------------------------------------------ planning
    def plan(self, read_lengths):
        cap = self.chunk_windows or self.m.max_chunk_windows()
        return JobPlan(read_lengths, self.T, self.step, self.start_index, self.batch, cap, self.m.chunk_align())

    def _buf(self, name, numel, dtype):
        t = self._bufs.get(name)
        if t is None or t.numel() < numel or t.dtype != dtype:
            self._bufs[name] = None
            t = torch.empty(max(int(numel), 1), dtype=dtype, device=self.m.device)
            self._bufs[name] = t
        return t

    def stage_raw(self, reads_or_raw, read_off=None):
        """Host staging of a rank's reads in PACKED order: returns (plan, pinned int16 buffer).  ``reads_or_raw`` is a
        list of int16 arrays, or one concatenated array with ``read_off``."""
        if read_off is None:
            reads = [np.asarray(r) for r in reads_or_raw]
            lengths = np.array([len(r) for r in reads], dtype=np.int64)
        else:
            read_off = np.asarray(read_off, dtype=np.int64)
            raw = np.asarray(reads_or_raw)
            reads = [raw[a:b] for a, b in zip(read_off[:-1], read_off[1:])]
            lengths = np.diff(read_off)
        plan = self.plan(lengths)
        host = torch.empty(max(plan.total_samples, 1), dtype=torch.int16, pin_memory=True).numpy()
        for pos, r in enumerate(plan.order):
            host[plan.read_off[pos]:plan.read_off[pos + 1]] = reads[r]
        return plan, host

    # ------------------------------------------------------------------ the device chain
    def run_device(self, raw_d, plan, keep_labels=False):
        """raw_d: int16 cuda tensor holding the reads of ``plan`` back to back in packed order.  Enqueues the whole
        chain on the current stream and returns (cons_d uint8 [capacity], cons_off_d int64 [n_reads + 1], status_d)."""
        m, T, dev = self.m, self.T, self.m.device
        n, nr = plan.n_windows, plan.n_reads
        st = _stream_ptr(dev)
        launches = 0
        with torch.cuda.device(dev):
            tab = self._buf("tab", 2 * (nr + 1), torch.int64)
            tab_h = np.concatenate([plan.read_off, plan.win_off])
            tab[:2 * (nr + 1)].copy_(torch.from_numpy(tab_h), non_blocking=True)
            x = self._buf("x", n * T, torch.float32)
            lens = self._buf("lens", n, torch.int32)
            labels = self._buf("labels", n * T, torch.int32)
            llen = self._buf("llen", n, torch.int32)
            cons = self._buf("cons", n * T, torch.uint8)
            cons_off = self._buf("cons_off", nr + 1, torch.int64)
            status = self._buf("status", 1, torch.int32)
            if n:
                _lib.check(self.lib.ntb_pack_windows(_ptr(raw_d), _ptr(tab), C.c_void_p(tab.data_ptr() + 8 * (nr + 1)), nr, n,
                                                     T, self.step, self.start_index, _ptr(x), _ptr(lens), st))
                launches += 2
                g = self._buf("goff", len(plan.group_off) + len(plan.chunks), torch.int32)
                rel, at, where = [], 0, []
                for a, b in plan.chunks:
                    rel.append(plan.group_off[a:b + 1] - plan.group_off[a])
                    where.append(at)
                    at += b - a + 1
                g[:at].copy_(torch.from_numpy(np.concatenate(rel).astype(np.int32)), non_blocking=True)
                # ONE workspace sized for the largest chunk (growing it chunk by chunk would keep two alive at once)
                ws = m._workspace(max(m.workspace_bytes(int(plan.group_off[b] - plan.group_off[a]), b - a)
                                      for a, b in plan.chunks))
                for (a, b), o in zip(plan.chunks, where):
                    w0, w1 = int(plan.group_off[a]), int(plan.group_off[b])
                    _lib.check(self.lib.ntb_model_forward(
                        m._h, C.c_void_p(x.data_ptr() + 4 * w0 * T), C.c_void_p(lens.data_ptr() + 4 * w0),
                        C.c_void_p(g.data_ptr() + 4 * o), w1 - w0, b - a, _ptr(ws), ws.numel(), C.c_void_p(0),
                        C.c_void_p(labels.data_ptr() + 4 * w0 * T), C.c_void_p(llen.data_ptr() + 4 * w0), st))
                    launches += m.last_launches()
            cap = n * T
            aws = self._buf("asm_ws", self.lib.ntb_assemble_workspace_bytes(n, nr, T, cap), torch.uint8)
            win_off_h = np.ascontiguousarray(plan.win_off, dtype=np.int64)
            _lib.check(self.lib.ntb_assemble_reads(_ptr(labels), win_off_h.ctypes.data_as(C.c_void_p), nr, n, T, cap,
                                                   _ptr(aws), aws.numel(), _ptr(cons), _ptr(cons_off), C.c_void_p(0),
                                                   _ptr(status), st))
            launches += 7
        self.last = {"launches": launches, "labels": labels[:n * T].view(n, T) if keep_labels else None,
                     "llen": llen[:n] if keep_labels else None, "lens": lens[:n] if keep_labels else None}
        return cons, cons_off, status

    def run_host(self, host_raw, plan, keep_labels=False):
        """End to end from a (pinned) host int16 buffer in packed order: H2D of the samples, the chain, D2H of the
        consensus.  Returns (cons uint8 numpy, cons_off int64 numpy [n_reads + 1]) in PACKED read order."""
        dev = self.m.device
        with torch.cuda.device(dev):
            raw_d = self._buf("raw", plan.total_samples, torch.int16)
            raw_d[:plan.total_samples].copy_(torch.from_numpy(host_raw[:plan.total_samples]), non_blocking=True)
            cons, cons_off, status = self.run_device(raw_d, plan, keep_labels=keep_labels)
            return self.fetch(cons, cons_off, status, plan.n_reads)

    def fetch(self, cons, cons_off, status, n_reads):
        """Device consensus -> host: offsets and status first (they size the second copy)."""
        off_h = torch.empty(n_reads + 2, dtype=torch.int64, pin_memory=True)
        off_h[:n_reads + 1].copy_(cons_off[:n_reads + 1], non_blocking=True)
        st_h = torch.empty(1, dtype=torch.int32, pin_memory=True)
        st_h.copy_(status, non_blocking=True)
        torch.cuda.current_stream(self.m.device).synchronize()
        if int(st_h[0]) != 0:
            raise _lib.NtbError(-1, "ntb_assemble_reads: status %d" % int(st_h[0]))
        total = int(off_h[n_reads])
        out = torch.empty(max(total, 1), dtype=torch.uint8, pin_memory=True)
        out[:total].copy_(cons[:total], non_blocking=True)
        torch.cuda.current_stream(self.m.device).synchronize()
        return out.numpy()[:total], off_h.numpy()[:n_reads + 1]

    @staticmethod
    def to_original_order(cons, cons_off, order):
        """Packed-order consensus -> list of byte strings in the caller's read order."""
        out = [None] * len(order)
        for pos, r in enumerate(order):
            out[int(r)] = cons[cons_off[pos]:cons_off[pos + 1]].tobytes()
        return out

    # ------------------------------------------------------------------ .signal files
    def run_signal_files(self, paths):
        """``.signal`` text files -> consensus strings (caller's order), parsed on the device (ntb_parse_signal)."""
        dev = self.m.device
        blobs = []
        for p in paths:
            with open(p, "rb") as fh:
                blobs.append(fh.read())
        file_off = np.zeros(len(blobs) + 1, dtype=np.int64)
        file_off[1:] = np.cumsum([len(b) for b in blobs])
        total = int(file_off[-1])
        text_h = _pin(np.frombuffer(b"".join(blobs), dtype=np.uint8).copy()) if total else np.zeros(1, np.uint8)
        capacity = total // 2 + len(blobs) + 1
        with torch.cuda.device(dev):
            st = _stream_ptr(dev)
            text = self._buf("text", total, torch.uint8)
            text[:total].copy_(torch.from_numpy(text_h[:total]), non_blocking=True)
            samples = self._buf("samples", capacity, torch.int16)
            roff = self._buf("parse_off", len(blobs) + 1, torch.int64)
            status = self._buf("parse_status", 1, torch.int32)
            pws = self._buf("parse_ws", self.lib.ntb_parse_signal_workspace_bytes(total, len(blobs)), torch.uint8)
            _lib.check(self.lib.ntb_parse_signal(_ptr(text), file_off.ctypes.data_as(C.c_void_p), len(blobs), total, _ptr(pws),
                                                 pws.numel(), capacity, _ptr(samples), _ptr(roff), _ptr(status), st))
            roff_h = roff[:len(blobs) + 1].cpu().numpy()           # sample counts size every later buffer
            if int(status.cpu()[0]) != 0:
                raise ValueError("ntb_parse_signal: status %d (2 = malformed token, 3 = value outside int16)" % int(status.cpu()[0]))
            lengths = np.diff(roff_h)
            if len(lengths) and lengths.min() == 0:
                raise ValueError("empty signal file %s" % paths[int(np.argmin(lengths))])
            plan = self.plan(lengths)
            # reorder the parsed reads into packed order on the device (one strided copy per read would be slow for
            # many reads; a gather index over samples keeps it one launch of torch's copy engine plumbing)
            if np.array_equal(plan.order, np.arange(len(lengths))):
                raw_d = samples
            else:
                raw_d = self._buf("raw", plan.total_samples, torch.int16)
                for pos, r in enumerate(plan.order):
                    raw_d[plan.read_off[pos]:plan.read_off[pos + 1]].copy_(samples[roff_h[r]:roff_h[r + 1]], non_blocking=True)
            cons, cons_off, status2 = self.run_device(raw_d, plan)
            cons_h, off_h = self.fetch(cons, cons_off, status2, plan.n_reads)
        return [s.decode("ascii") for s in self.to_original_order(cons_h, off_h, plan.order)]


def gather_consensus(cons_d, cons_off_d, n_reads, read_ids, n_total_reads, rank, world, device):
    """Brings every rank's called reads to rank 0 in ORIGINAL read order.  cons_d / cons_off_d: this rank's
    consensus bytes and offsets (packed order) on ``device``; read_ids [n_reads]: the original index of each packed
    read.  Sizes are exchanged with one small all_gather, the byte buffers with one padded gather (NCCL over NVLink
    when the buffers are in HBM; gloo in the CPU tests), then rank 0 makes ONE device->host copy and scatters by read id
    on the host.  Returns the list of bytes per original read on rank 0, None elsewhere."""
    import contextlib
    device = torch.device(device)
    cuda = device.type == "cuda"
    with (torch.cuda.device(device) if cuda else contextlib.nullcontext()):
        off_mine = cons_off_d[:n_reads + 1]
        meta = torch.stack([off_mine[n_reads], torch.tensor(n_reads, dtype=torch.int64, device=device)])
        if world == 1:
            metas = meta.view(1, 2).cpu().numpy()
        else:
            import torch.distributed as dist
            allm = torch.empty(world * 2, dtype=torch.int64, device=device)
            dist.all_gather_into_tensor(allm, meta)
            metas = allm.view(world, 2).cpu().numpy()               # sizes are needed on the host to size the gather
        max_bytes = int(metas[:, 0].max())
        max_reads = int(metas[:, 1].max())
        pad_b = -(-max(max_bytes, 1) // 16) * 16
        ids_d = torch.from_numpy(np.ascontiguousarray(read_ids, dtype=np.int64)).to(device, non_blocking=True)
        # one packed message per rank: [offsets (max_reads + 1) int64 | read ids (max_reads) int64 | bytes]
        hdr = torch.zeros(2 * max_reads + 1, dtype=torch.int64, device=device)
        hdr[:n_reads + 1] = off_mine
        hdr[max_reads + 1:max_reads + 1 + n_reads] = ids_d
        hb = hdr.numel() * 8
        msg = torch.empty(hb + pad_b, dtype=torch.uint8, device=device)
        msg[:hb] = hdr.view(torch.uint8)
        msg[hb:hb + int(metas[rank, 0])] = cons_d[:int(metas[rank, 0])]
        if world == 1:
            allmsg = msg.view(1, -1)
        else:
            allmsg = torch.empty((world, msg.numel()), dtype=torch.uint8, device=device) if rank == 0 else None
            dist.gather(msg, list(allmsg.unbind(0)) if rank == 0 else None, dst=0)
        if rank != 0:
            return None
        if cuda:
            host = torch.empty(allmsg.shape, dtype=torch.uint8, pin_memory=True)
            host.copy_(allmsg, non_blocking=True)
            torch.cuda.current_stream(device).synchronize()
        else:
            host = allmsg
        host = host.numpy()
    out = [None] * n_total_reads
    for r in range(world):
        h = host[r, :hb].view(np.int64)
        nr = int(metas[r, 1])
        off, ids = h[:nr + 1], h[max_reads + 1:max_reads + 1 + nr]
        body = host[r, hb:]
        for k in range(nr):
            out[int(ids[k])] = body[off[k]:off[k + 1]].tobytes()
    return out
