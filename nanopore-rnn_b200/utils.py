"""Config helpers used by the driver (restates nanotensor/utils.py:158-163 DotDict, :217-223
load_json, :44-66 list_dir, :150-155 merge_two_dicts of the reference)."""
import json
import os
import time


class DotDict(dict):
    """dict with attribute access; nested dicts are wrapped on read."""

    def __getattr__(self, name):
        try:
            v = self[name]
        except KeyError as e:
            raise AttributeError(name) from e
        if isinstance(v, dict) and not isinstance(v, DotDict):
            v = DotDict(v)
            self[name] = v
        return v

    __setattr__ = dict.__setitem__
    __delattr__ = dict.__delitem__


def load_json(path):
    with open(path) as fh:
        return json.load(fh)


def save_json(obj, path):
    with open(path, "w") as fh:
        json.dump(obj, fh, indent=2)
    return path


def list_dir(path, ext=""):
    """Absolute paths of the regular files in ``path`` whose last dot-separated component equals ``ext`` (all regular
    files when ``ext`` is empty): the reference's predicate, nanotensor/utils.py:71-82 (``isfile`` and
    ``f.split(".")[-1] == ext``).  Deliberate deviation: the result is sorted (the reference returns ``os.listdir``
    order, which is filesystem dependent) so that jobs, BN groups over concatenated files and shards are reproducible."""
    root = os.path.abspath(path)
    names = [f for f in os.listdir(path) if os.path.isfile(os.path.join(root, f))
             and (ext == "" or f.split(".")[-1] == ext)]
    return sorted(os.path.join(root, f) for f in names)


def merge_two_dicts(a, b):
    out = dict(a)
    out.update(b)
    return out


def time_it(fn, *args, **kw):
    t0 = time.perf_counter()
    fn(*args, **kw)
    return time.perf_counter() - t0
