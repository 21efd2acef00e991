"""CTC decoders on the GPU with the TensorFlow call signatures the reference uses:

* ``ctc_greedy_decoder(inputs, sequence_length, merge_repeated=True)`` -> (decoded, neg_sum_logits)
  (nanotensor/network.py:528-533, chiron/chiron/chiron_rcnn_train.py:68);
* ``ctc_beam_search_decoder(inputs, sequence_length, beam_width=100, top_paths=1, merge_repeated=True)`` ->
  (decoded, log_probabilities) (chiron/chiron/chiron_rcnn_train.py:66-67).

``inputs`` are time major ``[max_time, batch, num_classes]`` as in TensorFlow (``time_major=False`` takes the graphs'
``[batch, max_time, num_classes]`` output_layer directly); the blank is the last class.  ``decoded`` is a list of dense
int64 arrays ``[batch, longest]`` padded with -1 -- what ``tf.sparse_tensor_to_dense(decoded[k], default_value=-1)``
gives at the reference's call site.  Kernels: csrc/ntb_ctc.cuh behind ``ntb_ctc_greedy`` / ``ntb_ctc_beam_search``;
no CPU fallback."""
import ctypes as C

import numpy as np
import torch

from . import _lib


def _prep(inputs, sequence_length, time_major, device):
    x = torch.as_tensor(np.asarray(inputs, dtype=np.float32) if not isinstance(inputs, torch.Tensor) else inputs)
    if x.dim() != 3:
        raise ValueError("inputs must be [max_time, batch, num_classes]")
    if time_major:
        x = x.permute(1, 0, 2)
    dev = torch.device(device)
    if dev.type != "cuda" or not torch.cuda.is_available():
        raise _lib.NtbError(-2, "nanotensor_b200 needs a CUDA sm_100 device; there is no CPU fallback")
    x = x.to(dev, dtype=torch.float32).contiguous()
    lens = torch.as_tensor(np.asarray(sequence_length, dtype=np.int32)).to(dev)
    if lens.numel() != x.shape[0]:
        raise ValueError("sequence_length must hold one entry per batch element")
    return x, lens, dev


def _dense(rows, lens):
    width = int(lens.max()) if len(lens) else 0
    return rows[:, :width].astype(np.int64)


def ctc_greedy_decoder(inputs, sequence_length, merge_repeated=True, time_major=True, device="cuda:0"):
    x, lens, dev = _prep(inputs, sequence_length, time_major, device)
    n, T, Cn = x.shape
    lib = _lib.load()
    with torch.cuda.device(dev):
        labels = torch.empty((n, T), dtype=torch.int32, device=dev)
        llen = torch.empty(n, dtype=torch.int32, device=dev)
        neg = torch.empty(n, dtype=torch.float32, device=dev)
        _lib.check(lib.ntb_ctc_greedy(C.c_void_p(x.data_ptr()), C.c_void_p(lens.data_ptr()), n, T, Cn, int(bool(merge_repeated)),
                                      C.c_void_p(labels.data_ptr()), C.c_void_p(llen.data_ptr()), C.c_void_p(neg.data_ptr()),
                                      C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)))
        labels, llen, neg = labels.cpu().numpy(), llen.cpu().numpy(), neg.cpu().numpy()
    return [_dense(labels, llen)], neg.reshape(n, 1)


def ctc_beam_search_decoder(inputs, sequence_length, beam_width=100, top_paths=1, merge_repeated=True, time_major=True,
                            log_softmax=False, device="cuda:0"):
    x, lens, dev = _prep(inputs, sequence_length, time_major, device)
    n, T, Cn = x.shape
    lib = _lib.load()
    with torch.cuda.device(dev):
        ws = torch.empty(max(int(lib.ntb_ctc_beam_workspace_bytes(n, T, beam_width)), 1), dtype=torch.uint8, device=dev)
        paths = torch.empty((n, top_paths, T), dtype=torch.int32, device=dev)
        plen = torch.empty((n, top_paths), dtype=torch.int32, device=dev)
        logp = torch.empty((n, top_paths), dtype=torch.float32, device=dev)
        _lib.check(lib.ntb_ctc_beam_search(C.c_void_p(x.data_ptr()), C.c_void_p(lens.data_ptr()), n, T, Cn, int(beam_width),
                                           int(top_paths), int(bool(merge_repeated)), int(bool(log_softmax)),
                                           C.c_void_p(ws.data_ptr()), ws.numel(), C.c_void_p(paths.data_ptr()),
                                           C.c_void_p(plen.data_ptr()), C.c_void_p(logp.data_ptr()),
                                           C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)))
        paths, plen, logp = paths.cpu().numpy(), plen.cpu().numpy(), logp.cpu().numpy()
    return [_dense(paths[:, k], plen[:, k]) for k in range(top_paths)], logp
