"""nanopore-rnn_b200: B200-native forward of NanoTensor's BLSTM basecaller
(BD2KGenomics/nanopore-RNN, nanotensor/network.py as driven by run_nanotensor.py), a drop-in
for that path only.  Compute lives in libnanotensor_b200.so (hand-written sm_100a CUDA behind
the C ABI of include/nanotensor_b200.h); this package is the Python host side mirroring the
reference's classes.  No CPU fallback."""
from . import checkpoint, synthetic, utils  # noqa: F401
from .error import KernelLibraryMissing, PathError, Usage  # noqa: F401

__version__ = "0.1.0"


def __getattr__(name):
    # torch-dependent modules load lazily so that config / checkpoint tooling imports fast
    import importlib
    if name in ("engine", "network", "dataset", "run_nanotensor", "shard", "features", "assembly", "eventdetection", "_lib"):
        return importlib.import_module("." + name, __name__)
    raise AttributeError(name)
