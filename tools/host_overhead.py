"""Where the fixed host-side cost of engine.Model.forward_host goes (cProfile of a few calls on a 262 144-window job)."""
import cProfile
import os
import pstats
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import nanopore_rnn_b200  # noqa: E402,F401
from nanopore_rnn_b200 import engine  # noqa: E402

x, lens, goff = bench.make_workload(262144, 100)
net, w = bench.network_and_weights()
m = engine.Model(net, w, "CtcLoss", 100, 1, 4, device="cuda:0")
xp = torch.from_numpy(x).pin_memory().numpy()
lp = torch.from_numpy(lens).pin_memory().numpy()
for _ in range(2):
    m.forward_host(xp, lp, goff)
torch.cuda.synchronize()
t0 = time.perf_counter()
pr = cProfile.Profile()
pr.enable()
for _ in range(3):
    r = m.forward_host(xp, lp, goff)
pr.disable()
torch.cuda.synchronize()
print("ms per call", (time.perf_counter() - t0) / 3 * 1e3)
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
