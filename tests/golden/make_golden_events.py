#!/usr/bin/env python
"""Generates tests/golden/event_features.npz by RUNNING THE REFERENCE'S OWN CODE (run in the build container, where
/root/reference exists; the GPU box only reads the committed output).

SURVEY.md section 8(f) rank 4: the event-feature front end of the CrossEntropy (event) model,
``nanonet/nanonet/features.py::events_to_features`` (12 features per event: mean / stdv / length / delta at offsets
-1, 0, +1, each field standardised over the read).  That module cannot be imported as a whole (it pulls in
``nanonet.fast5`` -> h5py and a compiled event detector), but the functions on this path are plain NumPy: they are cut
out of the reference file with ``ast`` -- verbatim, by name -- and executed here, so the vectors below are the
reference's outputs, not a restatement's.  This row of the oracle is therefore PINNED.

Inputs are synthetic event tables (structured arrays with the fields the reference reads: mean, stdv, length), seeds and
shapes chosen to cover: ordinary reads, a 1-event and a 2-event read, a read whose stdv field is constant (scale 0 -> 1),
and a read with a large offset (the reference's second centring pass).
"""
import ast
import os

import numpy as np
import numpy.lib.recfunctions as nprf

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/nanonet/nanonet/features.py"
WANTED = ("padded_offset_array", "scale_array", "events_to_features", "SquiggleFeatureGenerator")


def reference_namespace():
    src = open(REF).read()
    tree = ast.parse(src)
    picked = [n for n in tree.body if isinstance(n, (ast.FunctionDef, ast.ClassDef)) and n.name in WANTED]
    assert sorted(n.name for n in picked) == sorted(WANTED), [n.name for n in picked]
    ns = {"np": np, "nprf": nprf}
    exec(compile(ast.Module(body=picked, type_ignores=[]), REF, "exec"), ns)
    return ns


def synthetic_events(n, seed, offset=0.0, const_stdv=False):
    rng = np.random.default_rng(seed)
    ev = np.zeros(n, dtype=[("start", "f8"), ("length", "f8"), ("mean", "f8"), ("stdv", "f8")])
    ev["length"] = rng.integers(3, 60, n) / 4000.0                     # dwell in seconds
    ev["start"] = np.concatenate([[0.0], np.cumsum(ev["length"])[:-1]])
    level = rng.choice(np.linspace(60.0, 120.0, 64), n)                # k-mer current levels, pA
    ev["mean"] = level + rng.normal(0.0, 1.5, n) + offset
    ev["stdv"] = 1.25 if const_stdv else np.abs(rng.normal(1.5, 0.4, n)) + 0.2
    return ev


def main():
    ns = reference_namespace()
    cases = [(1000, 1, 0.0, False), (1, 2, 0.0, False), (2, 3, 0.0, False), (257, 4, 0.0, True),
             (5000, 5, 1.0e9, False), (37, 6, 0.0, False)]
    fields, offs, outs = {"mean": [], "stdv": [], "length": []}, [0], []
    for n, seed, offset, const in cases:
        ev = synthetic_events(n, seed, offset, const)
        X = ns["events_to_features"](ev.copy())                        # the reference's own function
        assert X.shape == (n, 12) and X.dtype == np.float64
        for k in fields:
            fields[k].append(np.asarray(ev[k], dtype=np.float64))
        offs.append(offs[-1] + n)
        outs.append(X)
    np.savez_compressed(os.path.join(HERE, "event_features.npz"),
                        mean=np.concatenate(fields["mean"]), stdv=np.concatenate(fields["stdv"]),
                        length=np.concatenate(fields["length"]), read_off=np.asarray(offs, dtype=np.int64),
                        features=np.concatenate(outs))
    print("wrote event_features.npz: %d reads, %d events" % (len(cases), offs[-1]))


if __name__ == "__main__":
    main()
