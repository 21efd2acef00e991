"""Exceptions of the drop-in surface (mirrors nanotensor/error.py:25-40 of the reference)."""


class Usage(Exception):
    """Raised for a bad config / bad command line, as the reference driver does
    (run_nanotensor.py:567,575-576)."""

    def __init__(self, msg):
        super().__init__(msg)
        self.msg = msg


class PathError(Exception):
    """Raised for a missing file or directory."""

    def __init__(self, msg):
        super().__init__(msg)
        self.msg = msg


class KernelLibraryMissing(RuntimeError):
    """The sm_100a shared library is absent or cannot be loaded: there is no CPU fallback."""
