"""TEST INFRASTRUCTURE -- ctypes front end of oracle/oracle_cpu.cpp (C++/OpenMP restatement of the reference's
forward; parity unpinned, see the header of the .cpp).  Only tests/, smoke() and bench.py's CPU legs import this."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle_cpu.so")
_ORDER = ["conv1", "conv1_bn_scale", "conv1_bn_offset", "conv2a", "conv2a_bn_scale", "conv2a_bn_offset",
          "conv2b", "conv2b_bn_scale", "conv2b_bn_offset", "conv2c", "conv2c_bn_scale", "conv2c_bn_offset"]


class _Layer(C.Structure):          # include/nanotensor_b200.h: ntb_layer
    _fields_ = [("type", C.c_int32), ("in_dim", C.c_int32), ("size", C.c_int32), ("batchnorm", C.c_int32),
                ("forget_bias", C.c_float), ("w", C.POINTER(C.c_float) * 12)]


class _Desc(C.Structure):           # include/nanotensor_b200.h: ntb_model_desc
    _fields_ = [("graph", C.c_int32), ("seq_len", C.c_int32), ("len_x", C.c_int32), ("n_classes", C.c_int32),
                ("n_layers", C.c_int32), ("layers", C.POINTER(_Layer)), ("out_w", C.POINTER(C.c_float)),
                ("out_b", C.POINTER(C.c_float))]


_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("%s missing: run oracle/build.sh (or __graft_entry__.build())" % LIB_PATH)
        _lib = C.CDLL(LIB_PATH)
        _lib.oracle_cpu_forward.restype = C.c_int
        _lib.oracle_cpu_forward.argtypes = [C.POINTER(_Desc), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                            C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32]
        _lib.oracle_cpu_max_threads.restype = C.c_int
    return _lib


def max_threads():
    return int(load().oracle_cpu_max_threads())


def forward(x, lens, network, weights, graph, group_off=None, batch_size=0, n_threads=0):
    """-> (logits [n*T, C] float32, labels, label_len).  labels: CrossEntropy [n*T] int32; CtcLoss [n, T] int32
    padded with -1.  group_off: window offsets of the reference's batches (BN groups)."""
    lib = load()
    x = np.ascontiguousarray(x, dtype=np.float32)
    lens = np.ascontiguousarray(lens, dtype=np.int32)
    n, T = x.shape[0], x.shape[1]
    len_x = 1 if x.ndim == 2 else x.shape[2]
    keep = []

    def fp(a):
        a = np.ascontiguousarray(a, dtype=np.float32)
        keep.append(a)
        return a.ctypes.data_as(C.POINTER(C.c_float))

    layers = (_Layer * len(network))()
    dim = len_x
    for L, l, w in zip(layers, network, weights["layers"]):
        L.in_dim = dim
        if l["type"] == "residual_layer":
            L.type, L.size, L.batchnorm = 1, int(l["out_channel"]), int(bool(l.get("batchnorm")))
            for k, name in enumerate(_ORDER):
                if name in w:
                    L.w[k] = fp(w[name])
            dim = L.size
        elif l["type"] == "blstm":
            L.type, L.size, L.forget_bias = 2, int(l["size"]), float(l["bias"])
            for k, name in enumerate(["kernel_fw", "bias_fw", "kernel_bw", "bias_bw"]):
                L.w[k] = fp(w[name])
            dim = 2 * L.size
        elif l["type"] == "chiron_fnn":
            L.type, L.size = 3, dim // 2
            L.w[0], L.w[1] = fp(w["weights"]), fp(w["bias"])
            dim //= 2
        else:
            raise ValueError(l["type"])
    Cn = int(weights["out_w"].shape[1])
    desc = _Desc(1 if graph == "CtcLoss" else 0, T, len_x, Cn, len(network), layers, fp(weights["out_w"]),
                 fp(weights["out_b"]))
    logits = np.empty((n * T, Cn), dtype=np.float32)
    labels = np.empty(n * T, dtype=np.int32)
    llen = np.zeros(n, dtype=np.int32)
    go, ng = None, 0
    if group_off is not None:
        go = np.ascontiguousarray(group_off, dtype=np.int32)
        ng = len(go) - 1
    rc = lib.oracle_cpu_forward(C.byref(desc), x.ctypes.data, lens.ctypes.data, go.ctypes.data if go is not None else None,
                                n, ng, int(batch_size), logits.ctypes.data, labels.ctypes.data, llen.ctypes.data,
                                int(n_threads))
    if rc != 0:
        raise RuntimeError("oracle_cpu_forward failed (%d)" % rc)
    if graph == "CtcLoss":
        return logits, labels.reshape(n, T), llen
    return logits, labels, None
