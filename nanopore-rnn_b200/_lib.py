"""ctypes binding of libnanotensor_b200.so (the C ABI in include/nanotensor_b200.h).

There is no CPU fallback: if the shared library is missing or cannot be loaded,
``KernelLibraryMissing`` is raised the first time a compute entry point is needed."""
import ctypes as C
import os

from .error import KernelLibraryMissing

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libnanotensor_b200.so")

NTB_LAYER_RESIDUAL, NTB_LAYER_BLSTM, NTB_LAYER_CHIRON_FNN = 1, 2, 3
NTB_GRAPH_CROSS_ENTROPY, NTB_GRAPH_CTC = 0, 1
NTB_MAX_LAYER_TENSORS = 12
_FP = C.POINTER(C.c_float)


class NtbLayer(C.Structure):
    _fields_ = [("type", C.c_int32), ("in_dim", C.c_int32), ("size", C.c_int32),
                ("batchnorm", C.c_int32), ("forget_bias", C.c_float),
                ("w", _FP * NTB_MAX_LAYER_TENSORS)]


class NtbModelDesc(C.Structure):
    _fields_ = [("graph", C.c_int32), ("seq_len", C.c_int32), ("len_x", C.c_int32),
                ("n_classes", C.c_int32), ("n_layers", C.c_int32),
                ("layers", C.POINTER(NtbLayer)), ("out_w", _FP), ("out_b", _FP)]


_P, _I64, _I32, _F = C.c_void_p, C.c_int64, C.c_int32, C.c_float
# name -> (restype, argtypes); must list every symbol include/nanotensor_b200.h declares
SIGNATURES = {
    "ntb_last_error": (C.c_char_p, []),
    "ntb_version": (C.c_char_p, []),
    "ntb_device_ok": (C.c_int, []),
    "ntb_model_weight_bytes": (_I64, [C.POINTER(NtbModelDesc)]),
    "ntb_model_create": (C.c_int, [C.POINTER(NtbModelDesc), _P, _I64, _P, C.POINTER(_P)]),
    "ntb_model_destroy": (None, [_P]),
    "ntb_model_workspace_bytes": (_I64, [_P, _I64, _I64]),
    "ntb_model_forward": (C.c_int, [_P, _P, _P, _P, _I64, _I64, _P, _I64, _P, _P, _P, _P]),
    "ntb_model_forward_host": (C.c_int, [_P, _P, _P, _P, _I64, _I64, _P, _I64, _P, _P, _P, _P]),
    "ntb_model_forward_host_async": (C.c_int, [_P, _P, _P, _P, _I64, _I64, _P, _I64, _P, _P, _P, _P]),
    "ntb_model_workspace_bytes_host": (_I64, [_P, _I64, _I64, C.c_int]),
    "ntb_model_last_launches": (_I64, [_P]),
    "ntb_model_set_profile": (C.c_int, [_P, C.c_int]),
    "ntb_model_stage_ms": (C.c_int, [_P, _FP, C.c_int]),
    "ntb_pack_windows": (C.c_int, [_P, _P, _P, _I64, _I64, _I32, _I32, _I32, _P, _P, _P]),
    "ntb_event_features_workspace_bytes": (_I64, [_I64, _I64]),
    "ntb_event_features": (C.c_int, [_P, _P, _P, _P, _I64, _I64, _I32, _P, _I64, _P, _P]),
    "ntb_event_features_host": (C.c_int, [_P, _P, _P, _P, _I64, _I64, _I32, _P, _P]),
    "ntb_assemble_workspace_bytes": (_I64, [_I64, _I64, _I32, _I64]),
    "ntb_assemble_reads": (C.c_int, [_P, _P, _I64, _I64, _I32, _I64, _P, _I64, _P, _P, _P, _P, _P]),
    "ntb_assemble_reads_host": (C.c_int, [_P, _P, _I64, _I64, _I32, _I64, _P, _P, _P, _P]),
    "ntb_event_detect_workspace_bytes": (_I64, [_I64, _I64]),
    "ntb_event_detect": (C.c_int, [_P, _P, _I64, _I64, C.c_double, _P, _P, C.c_double, _P, _I64, _I64, _P, _P, _P, _P]),
    "ntb_event_detect_host": (C.c_int, [_P, _P, _I64, _I64, C.c_double, _P, _P, C.c_double, _I64, _P, _P, _P]),
    "ntb_parse_signal_workspace_bytes": (_I64, [_I64, _I64]),
    "ntb_parse_signal": (C.c_int, [_P, _P, _I64, _I64, _P, _I64, _I64, _P, _P, _P, _P]),
    "ntb_parse_signal_host": (C.c_int, [_P, _P, _I64, _I64, _I64, _P, _P, _P]),
    "ntb_conv_bn": (C.c_int, [_P, _P, _P, _P, _P, _I64, _I64, _I32, _I32, _I32, _I32, _I32, _P, _I64, _P, _P]),
    "ntb_conv_bn_workspace_bytes": (_I64, [_I64, _I64, _I32, _I32, _I32, _I32]),
    "ntb_blstm": (C.c_int, [_P, _P, _P, _P, _P, _P, _F, _I64, _I32, _I32, _I32, _P, _I64, _P, _P]),
    "ntb_blstm_workspace_bytes": (_I64, [_I64, _I32, _I32, _I32]),
    "ntb_decode": (C.c_int, [_P, _P, _I64, _I32, _I32, _I32, _P, _P, _P]),
    "ntb_ctc_greedy": (C.c_int, [_P, _P, _I64, _I32, _I32, _I32, _P, _P, _P, _P]),
    "ntb_ctc_beam_workspace_bytes": (_I64, [_I64, _I32, _I32]),
    "ntb_ctc_beam_search": (C.c_int, [_P, _P, _I64, _I32, _I32, _I32, _I32, _I32, _I32, _P, _I64, _P, _P, _P, _P]),
    "ntb_ctc_beam_search_noalloc": (C.c_int, [_P, _P, _I64, _I32, _I32, _I32, _I32, _I32, _I32, _P, _P, _P, _P]),
    "ntb_gemm": (C.c_int, [_P, _P, _P, _I64, _I32, _I32, _I32, _P, _I64, _P, _P]),
    "ntb_gemm_workspace_bytes": (_I64, [_I64, _I32, _I32, _I32]),
}

_lib = None


def load():
    """Load the shared library once; raises KernelLibraryMissing when it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise KernelLibraryMissing(
            "%s not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU fallback." % LIB_PATH)
    try:
        lib = C.CDLL(LIB_PATH)
    except OSError as e:
        raise KernelLibraryMissing("cannot load %s: %s" % (LIB_PATH, e)) from e
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:
            raise KernelLibraryMissing("%s does not export %s" % (LIB_PATH, name)) from e
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class NtbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("nanotensor_b200 error %d: %s" % (code, msg))
        self.code = code


def check(rc):
    if rc != 0:
        raise NtbError(rc, load().ntb_last_error().decode("utf-8", "replace"))
    return rc
